"""Shared inference machinery of the accelerated models: weight-pack cache, program cache,
the frame-side (post length-regulator) program common to ForwardTacotron and
MultiForwardTacotron, and the host-side result assembly."""
from collections import OrderedDict
from typing import Dict, Optional, Tuple  # noqa: F401

import torch
import torch.nn as nn

from .. import engine as eng
from .. import pack as pk
from .. import runtime as rt


def pack_bn_conv(m, split: bool = False) -> pk.PackedGemm:
    bn = m.bnorm
    return pk.pack_bn_conv(m.conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var, m.relu is True,
                           split=split)


def pack_cbhg(m, split_bank: bool = False) -> eng.CbhgPack:
    """split_bank: the bank / proj1 / proj2 weights also keep their unrounded copy, so the conv chain can run on
    [hi | lo | hi] split operands (the packed-sequence frame program, whose input carries |value| ~ 30 padding rows)."""
    bank = [pack_bn_conv(c, split=split_bank) for c in m.conv1d_bank]
    hws = [pk.pack_highway(h.W1.weight, h.W1.bias, h.W2.weight, h.W2.bias) for h in m.highways]
    return eng.CbhgPack(bank, pack_bn_conv(m.conv_project1, split=split_bank), pack_bn_conv(m.conv_project2, split=split_bank),
                        pk.pack_linear(m.pre_highway.weight), hws, pk.PackedRnn(m.rnn),
                        m.conv1d_bank[0].conv.in_channels, m.conv1d_bank[0].conv.out_channels)


def pack_predictor(m, precise: bool = False) -> eng.PredictorPack:
    """precise: the split-precision form of the whole predictor (conv stack, W_ih projection, recurrence) -- the
    DURATION predictor, whose output is rounded to integers (forward_tacotron.py:174-177, common_layers.py:18-21)."""
    precise = precise and eng.PRECISE_DUR and m.rnn.hidden_size <= 128 and isinstance(m.rnn, torch.nn.GRU)
    return eng.PredictorPack([pack_bn_conv(c, split=precise) for c in m.convs], pk.PackedRnn(m.rnn, precise=precise),
                             m.lin.weight, m.lin.bias, precise=precise)


class AcceleratedMixin:
    """Caches derived state on the module; everything is rebuilt if a parameter changes."""

    MAX_PROGRAMS = 4

    def _accel_signature(self) -> Tuple:
        sig = [self.training]
        for t in list(self.parameters()) + list(self.buffers()):
            sig.append((t.data_ptr(), t._version, t.device))
        return tuple(sig)

    def _accel_state(self, device: torch.device) -> dict:
        sig = self._accel_signature()
        st = self.__dict__.get('_accel')
        if st is None or st['sig'] != sig or st['device'] != device:
            if st is not None:
                for prog in list(st['programs'].values()):
                    prog.close()
            # Some packs alias the parameters themselves (embedding tables, predictor heads, the pitch / energy
            # projections) and the kernels read them with 16-byte vector loads: a parameter that is a view at an odd
            # offset of a larger buffer would be read from the wrong address.  torch allocations are 512-byte aligned.
            for name, t in list(self.named_parameters()) + list(self.named_buffers()):
                if t.is_floating_point() and t.data_ptr() % 16:
                    raise RuntimeError(f'parameter {name!r} is not 16-byte aligned (a view into a larger buffer?): '
                                       'clone it or lay the buffer out in 16-byte steps')
            handle = rt.Handle.get(device)
            st = {'sig': sig, 'device': device, 'handle': handle, 'packs': self._build_packs(),
                  'programs': OrderedDict()}
            self.__dict__['_accel'] = st
        return st

    def _program(self, st: dict, key, builder):
        """key = (kind, batch, seq_len).  Programs are cached per CAPACITY bucket of the sequence length
        (engine.capacity_bucket) and re-pointed at the actual length (Program.set_seq_len: a few small uploads and
        strided memsets): the regulated length M is data dependent and new on almost every call with real text
        (common_layers.py:23), and a program owns ~20 buffers, tensor maps and a captured graph."""
        kind, batch, seq_len = key
        ckey = (kind, batch, eng.capacity_bucket(seq_len))
        progs = st['programs']
        if ckey in progs:
            progs.move_to_end(ckey)
            prog = progs[ckey]
        else:
            prog = builder(ckey[2])
            progs[ckey] = prog
            while len(progs) > self.MAX_PROGRAMS * 2:
                _, old = progs.popitem(last=False)
                # frame programs read the x-projection buffer of the token program they were built for
                for dep in getattr(old, 'dependents', ()):
                    d = progs.pop(dep, None)
                    if d is not None:
                        d.close()
                old.close()
        prog.set_seq_len(seq_len)
        return prog

    def program_for(self, kind: str, batch: int, seq_len: int):
        """The cached program that served (kind, batch, seq_len), or None (tests, bench.py)."""
        st = self.__dict__.get('_accel')
        if st is None:
            return None
        bucket = eng.capacity_bucket(seq_len)
        prog = st['programs'].get((kind, batch, bucket))
        if prog is None:            # frame programs are keyed '<kind>@<token program>' (frame_program below): latest match
            for (k, b, c), p in reversed(list(st['programs'].items())):
                if k.startswith(kind + '@') and b == batch and c == bucket:
                    return p
        return prog

    @staticmethod
    def _check_tokens(x: torch.Tensor):
        if not x.is_cuda:
            raise RuntimeError('generate(): x must be a CUDA tensor -- this package has no CPU path '
                               '(the sm_100a kernels are the implementation)')
        if x.dim() != 2 or x.dtype != torch.int64:
            raise RuntimeError(f'generate(): x must be int64 [batch, tokens], got {x.dtype} {tuple(x.shape)}')


def token_x_supported(packs: dict, in_channels: int) -> bool:
    """The BiLSTM's x-projection once per PHONEME (engine.XPROJ_TOKENS): the fp16 operand / fp16 x-projection path."""
    lstm = packs['lstm']
    # (the indexed x read of the recurrence kernel is instantiated for the LSTM of 512 units, rnn_dims of both yaml configs)
    return bool(eng.XPROJ_TOKENS and eng.F16_OPERANDS and eng.X16 and in_channels % 8 == 0 and lstm.ih.n % 8 == 0
                and lstm.cell == 1 and lstm.hidden == 512)


def token_x_program(handle: rt.Handle, packs: dict, tok_prog: eng.Program, in_channels: int) -> eng.Program:
    """The token program's companion (built on first use, same capacity): 'h_tok' = prenet output + pitch / energy
    projection adds (+ speaker tail) per PHONEME as the fp16 operand, and the W_ih half of self.lstm over it
    (forward_tacotron.py:214-224 with the LengthRegulator hoisted behind the projection: W_ih x + b_ih is a per-row map,
    so it commutes with the regulator's row repetition, common_layers.py:21).  Row cap_rows of 'xproj' holds b_ih, the
    projection of the regulator's zero padding (common_layers.py:23)."""
    xp = getattr(tok_prog, 'xprog', None)
    if xp is None:
        geo = eng.Geometry(tok_prog.geo.batch, tok_prog.geo.cap)
        xp = eng.Program(handle, geo)
        lstm: pk.PackedRnn = packs['lstm']
        h_tok = xp.buf('h_tok', in_channels, torch.float16)
        xproj = xp.dense('xproj', (geo.cap_rows + 1, lstm.ih.n), torch.float16)
        xproj[geo.cap_rows] = lstm.ih.bias0.to(torch.float16)
        xp.label('lstm_ih')
        xp.gemm(h_tok, lstm.ih, xproj, round_out=False)
        xp.flush()
        xp.label('')
        xp.pad_row = geo.cap_rows
        xp.ident = {}
        tok_prog.xprog = xp
        tok_prog.dependents = set()
    xp.set_seq_len(tok_prog.geo.seq_len)
    return xp


def token_xproj(handle: rt.Handle, packs: dict, tok_prog: eng.Program, pre: torch.Tensor, extra: Optional[torch.Tensor],
                c2: int, pitch_hat: torch.Tensor, energy_hat: torch.Tensor, pitch_strength: float,
                energy_strength: float) -> None:
    """Enqueue the per-phoneme x-projection of the BiLSTM.  Nothing here depends on the regulated length, so generate()
    calls it BEFORE it waits for the LengthRegulator scan: the GEMM runs while the scan result travels to the host."""
    B, T = tok_prog.geo.batch, tok_prog.geo.seq_len
    C1 = pre.shape[1]
    xp = token_x_program(handle, packs, tok_prog, C1 + c2)
    ident = xp.ident.get(T)
    if ident is None:                                  # the identity regulator: one frame per phoneme
        if len(xp.ident) > 4:
            xp.ident.clear()
        cum = torch.arange(1, T + 1, dtype=torch.int32, device=pre.device).expand(B, T).contiguous()
        ident = xp.ident[T] = (cum, torch.full((B,), T, dtype=torch.int32, device=pre.device))
    g = _gather_args(packs, tok_prog, pre, extra, c2, pitch_hat, energy_hat, pitch_strength, energy_strength)
    g.cum, g.lens = ident[0].data_ptr(), ident[1].data_ptr()
    g.batch, g.T, g.M = B, T, T
    h_tok = xp.bufs['h_tok']
    g.out, g.out_stride, g.out_ld = h_tok.data_ptr(), xp.geo.stride, h_tok.shape[1]
    g.idx, g.round_tf32 = None, 2
    g._keep = g._keep + [ident]
    handle.lr_gather(g)
    xp.run()
    tok_prog.x_token = (pitch_hat, energy_hat)


def frame_program(model, st: dict, kind: str, tok_prog: eng.Program, batch: int, M: int, in_channels: int, n_mels: int,
                  packed_pad_value: Optional[float] = None) -> eng.Program:
    """The cached frame-side program for (kind, batch, M).  With the per-phoneme x-projection its BiLSTM reads the
    'xproj' buffer of `tok_prog`'s companion program, so it is keyed by that token program too and leaves the cache
    with it (AcceleratedMixin._program)."""
    packs, h = st['packs'], st['handle']
    if not token_x_supported(packs, in_channels):
        return model._program(st, (kind, batch, M), lambda M: build_frame_program(h, packs, batch, M, in_channels, n_mels,
                                                                                  packed_pad_value))
    xp = token_x_program(h, packs, tok_prog, in_channels)
    key_kind = f'{kind}@{id(tok_prog)}'
    prog = model._program(st, (key_kind, batch, M), lambda M: build_frame_program(h, packs, batch, M, in_channels, n_mels,
                                                                                  packed_pad_value, xprog=xp))
    tok_prog.dependents.add((key_kind, batch, eng.capacity_bucket(M)))
    return prog


def build_frame_program(handle: rt.Handle, packs: dict, batch: int, M: int, in_channels: int, n_mels: int,
                        packed_pad_value: Optional[float] = None, xprog: Optional[eng.Program] = None):
    """LR output [B,M,C] -> BiLSTM -> lin -> CBHG postnet -> post_proj
    (forward_tacotron.py:224-231).  Returns the Program; the LR gather writes prog.buf('lr') -- or, with `xprog` (the
    per-phoneme x-projection, token_x_program), nothing is gathered at all: the BiLSTM reads xprog's 'xproj' rows through
    the frame -> row map the caller writes into prog.bufs['x_idx'] (ftb_lr_index).
    `packed_pad_value` != None: the BiLSTM runs with pack_padded_sequence semantics over the int32 lengths the
    caller writes into prog.bufs['mel_len'] (eval-mode forward(batch), forward_tacotron.py:147-152)."""
    geo = eng.Geometry(batch, M)
    prog = eng.Program(handle, geo)
    lstm: pk.PackedRnn = packs['lstm']
    x16 = eng.X16 and eng.F16_OPERANDS and lstm.ih.n % 8 == 0
    if xprog is not None:
        prog.dense('x_idx', (batch, -1), torch.int32)
        xmap = dict.__getitem__(prog.bufs, 'x_idx')                     # the storage: [batch, M] dense at its head
        xp = dict.__getitem__(xprog.bufs, 'xproj')
        prog._keep.append(xprog)
        x = None
    else:
        xmap = None
        # the regulated sequence feeds only the W_ih GEMM: fp16 operand copy (K5 writes it directly)
        x = prog.buf('lr', in_channels, torch.float16 if (eng.F16_OPERANDS and in_channels % 8 == 0) else torch.float32)
        xp = prog.buf('lstm.xproj', lstm.ih.n, torch.float16 if x16 else torch.float32)
    # the recurrences hand h_t on in fp16, so their layer outputs are fp16-exact: `lin` and `post_proj` (their only
    # readers) take them as fp16 operands
    odt = torch.float16 if eng.F16_OPERANDS else torch.float32
    hl = prog.buf('lstm.h', 2 * lstm.hidden, odt)
    # channels-last mel: only the postnet bank's operand (the residual uses mel_cl_full): fp16 copy
    mel_cl = prog.buf('mel_cl', n_mels, torch.float16 if (eng.F16_OPERANDS and n_mels % 8 == 0) else torch.float32,
                      lead=rt.FOLD_LEAD_ROWS)       # zero rows in front: the postnet bank folds its taps into K
    mel = prog.dense('mel', (batch, n_mels, -1))
    mel_post = prog.dense('mel_post', (batch, n_mels, -1))
    post = prog.buf('postnet.h', 2 * packs['postnet'].rnn.hidden, odt)

    if x is not None:
        prog.label('lstm_ih')
        prog.gemm(x, lstm.ih, xp, round_out=False)                # W_ih half of self.lstm (:224)
    prog.label('lstm')
    if packed_pad_value is None:
        prog.rnn([(lstm, xp, hl, 0, xmap)])
    else:
        mel_len = prog.dense('mel_len', (batch,), torch.int32)
        # the padded rows feed the `lin` GEMM: keep the pad value tf32-exact like every other operand
        prog.rnn([(lstm, xp, hl, 0, xmap)], lens=mel_len, pad_value=float(pk.round_tf32(torch.tensor([packed_pad_value]))[0]))
    prog.label('lin')
    mel_full = prog.buf('mel_cl_full', n_mels)
    prog.gemm(hl, packs['lin'], mel_cl, out_t=mel, out_full=mel_full)   # self.lin + transpose (:226-227)
    prog.flush()
    if packed_pad_value is not None:
        # Rows behind a packed sentence hold lin(padding vector): 1024 identical operands, so the tf32 rounding of
        # the pad value is a SYSTEMATIC error (2.4e-4 relative on values of magnitude ~30) that the postnet's bank
        # and BiGRU carry into the valid frames.  Overwrite those rows with the exact fp32 vector instead.
        g = prog.geo
        v = packs['lin_pad_vec']
        v_r = pk.round_tf32(v)
        t_all = torch.arange(g.stride, device=mel_cl.device, dtype=torch.int32)

        def fix_pad_rows():
            L = g.seq_len                                                            # current geometry of the program
            t_idx = t_all[:g.stride]
            m = ((t_idx[None, :] >= mel_len[:, None]) & (t_idx[None, :] < L))        # [B, stride]
            cl = mel_cl[:g.rows].view(batch, g.stride, n_mels)
            clf = mel_full[:g.rows].view(batch, g.stride, n_mels)
            torch.where(m[:, :, None], v_r.to(cl.dtype), cl, out=cl)
            torch.where(m[:, :, None], v, clf, out=clf)
            mel_now = prog.bufs['mel']
            torch.where(m[:, None, :L], v[None, :, None], mel_now, out=mel_now)
            if mel_split is not None:
                # The padding rows (|value| ~ 30) sit next to every packed sentence: a single fp16 / tf32 rounding of
                # operands of that size is a 7e-3 absolute error that the bank taps and the backward GRU carry into the
                # valid frames.  The postnet bank therefore reads the mel as a split operand [hi | lo | hi].
                hi = clf.to(torch.float16)
                sp = mel_split[:g.rows].view(batch, g.stride, 3 * seg)
                sp[:, :, :n_mels] = hi
                sp[:, :, seg:seg + n_mels] = (clf - hi.float()).to(torch.float16)
                sp[:, :, 2 * seg:2 * seg + n_mels] = hi
                sp[:, L:] = 0
        seg = pk.ceil_to(n_mels, 8)
        mel_split = prog.buf('mel_split', 3 * seg, torch.float16) if packs['postnet'].bank[0].w_src is not None else None
        prog.call(fix_pad_rows, label='lin.pad_rows', torch_ops=True)
    else:
        mel_split = None
    prog.label('')
    levels, gru = eng.cbhg_levels(prog, 'postnet', packs['postnet'], mel_cl, post, x_full=mel_full, xproj_f16=x16,
                                  x_split=mel_split)
    eng.run_levels(prog, [levels])
    prog.label('postnet.gru')
    prog.rnn([gru])
    prog.label('post_proj')
    prog.gemm(post, packs['post_proj'], None, out_t=mel_post)     # self.post_proj + transpose (:230-231)
    prog.flush()
    return prog


# Set by forwardtacotron_b200.parallel around a sharded generate(): called with (meta, lens) -- both device tensors --
# right after K5's scan has been enqueued and BEFORE this rank's host blocks on the scan result, so a collective it
# enqueues (the exchange of the regulated lengths between the ranks) runs on the GPU while the host is waiting anyway.
scan_hook = None
_PINNED_META = {}


class PendingScan:
    """K5's scan enqueued, its 16-byte result on the way to pinned host memory: `finish()` waits for that copy."""

    def __init__(self, cum, lens, host, event):
        self.cum, self.lens, self.host, self.event = cum, lens, host, event

    def finish(self):
        self.event.synchronize()
        host = self.host
        if int(host[3]) != 0:
            # what nn.Embedding raises in the reference (forward_tacotron.py:31,210) for an id outside its table; the
            # kernels clamped the id instead of reading out of bounds, the outputs of this call are meaningless
            raise IndexError('index out of range in self: a token / class id lies outside its embedding table')
        return self.cum, self.lens, int(host[0]), int(host[2])


def length_regulate_begin(handle: rt.Handle, dur: torch.Tensor, apply_fallback: bool) -> PendingScan:
    """Enqueues K5's scan on `dur` (clamped / fallback-filled IN PLACE like the reference) and the copy of the three
    integers the host needs (max length, fallback flag, total frames) into a pinned buffer, followed by an event: the one
    D2H of generate().  The caller keeps enqueueing work that does not depend on the regulated length (the pitch / energy
    callbacks) and calls `finish()` when it needs M -- the host then waits for the 16-byte copy, not for a stream drain."""
    B, T = dur.shape
    dev = dur.device
    cum = torch.empty((B, T), dtype=torch.int32, device=dev)
    lens = torch.empty((B,), dtype=torch.int32, device=dev)
    meta = torch.empty((4,), dtype=torch.int32, device=dev)
    handle.lr_scan(dur, B, T, apply_fallback, cum, lens, meta)
    pin = _PINNED_META.get(dev)
    if pin is None:
        pin = _PINNED_META[dev] = (torch.empty((4,), dtype=torch.int32).pin_memory(), torch.cuda.Event())
    host, ev = pin
    host.copy_(meta, non_blocking=True)
    ev.record(torch.cuda.current_stream(dev))
    if scan_hook is not None:
        scan_hook(meta, lens)
    return PendingScan(cum, lens, host, ev)


def length_regulate(handle: rt.Handle, dur: torch.Tensor, apply_fallback: bool):
    """length_regulate_begin(...).finish(): (cum, lens, M, total frames)."""
    return length_regulate_begin(handle, dur, apply_fallback).finish()


def _gather_args(packs: dict, tok_prog: eng.Program, pre: torch.Tensor, extra: Optional[torch.Tensor], c2: int,
                 pitch_hat: torch.Tensor, energy_hat: torch.Tensor, pitch_strength: float,
                 energy_strength: float) -> rt.LrGatherArgs:
    """Source side of K5/K6's gather: prenet rows, speaker tail, the pitch / energy projection adds."""
    B, T = tok_prog.geo.batch, tok_prog.geo.seq_len
    pitch = pitch_hat.reshape(B, T).to(torch.float32).contiguous()
    energy = energy_hat.reshape(B, T).to(torch.float32).contiguous()
    C1 = pre.shape[1]
    g = rt.LrGatherArgs()
    g.x, g.x_stride, g.x_ld, g.c1 = pre.data_ptr(), tok_prog.geo.stride, C1, C1
    g.extra, g.c2 = rt.ptr(extra) if c2 else None, c2
    g.pitch, g.energy = pitch.data_ptr(), energy.data_ptr()
    (wp, bp), (we, be) = packs['pitch_proj'], packs['energy_proj']
    g.wp, g.bp, g.ps = wp.data_ptr(), bp.data_ptr(), pitch_strength
    g.we, g.be, g.es = we.data_ptr(), be.data_ptr(), energy_strength
    g._keep = [pitch, energy]
    return g


def gather_and_run_frames(handle: rt.Handle, packs: dict, tok_prog: eng.Program, prog: eng.Program,
                          pre: torch.Tensor, extra: Optional[torch.Tensor], c2: int, dur_hat: torch.Tensor,
                          pitch_hat: torch.Tensor, energy_hat: torch.Tensor, cum: torch.Tensor, lens: torch.Tensor,
                          M: int, pitch_strength: float, energy_strength: float,
                          round_tf32: bool = True) -> Dict[str, torch.Tensor]:
    """K5/K6 gather (projection adds + optional speaker tail fused into the source read) into the frame
    program's 'lr' buffer, then the frame-side launches; returns the reference's result dict
    (forward_tacotron.py:214-233).  Frame programs built on the per-phoneme x-projection (build_frame_program(xprog=))
    get the frame -> row map instead of the gathered rows."""
    B, T = dur_hat.shape
    if 'x_idx' in prog.bufs:
        xt = getattr(tok_prog, 'x_token', None)
        if xt is None or xt[0] is not pitch_hat or xt[1] is not energy_hat:       # not enqueued ahead of the scan wait
            token_xproj(handle, packs, tok_prog, pre, extra, c2, pitch_hat, energy_hat, pitch_strength, energy_strength)
        tok_prog.x_token = None
        xp = tok_prog.xprog
        handle.lr_index(cum, lens, B, T, M, prog.bufs['x_idx'], row_stride=xp.geo.stride, pad_row=xp.pad_row)
    else:
        g = _gather_args(packs, tok_prog, pre, extra, c2, pitch_hat, energy_hat, pitch_strength, energy_strength)
        g.cum, g.lens = cum.data_ptr(), lens.data_ptr()
        g.batch, g.T, g.M = B, T, M
        lr = prog.bufs['lr']
        g.out, g.out_stride, g.out_ld = lr.data_ptr(), prog.geo.stride, lr.shape[1]
        g.idx, g.round_tf32 = None, (2 if lr.dtype == torch.float16 else int(round_tf32))
        handle.lr_gather(g)
    prog.run()
    mel = prog.bufs['mel'].clone()
    mel_post = prog.bufs['mel_post'].clone() if 'mel_post' in prog.bufs else mel
    return {'mel': mel, 'mel_post': mel_post, 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}


# ------------------------------------------------------------------------------------------------
# FFT-block transformer (FastPitch / MultiFastPitch; common_layers.py:148-223)
# ------------------------------------------------------------------------------------------------

class FftLayerPack:
    def __init__(self, layer, split: bool = False):
        a = layer.self_attn
        self.in_proj = pk.pack_linear(a.in_proj_weight, a.in_proj_bias, split=split)
        self.out_proj = pk.pack_linear(a.out_proj.weight, a.out_proj.bias, split=split)
        self.conv1 = pk.pack_conv(layer.conv1.weight, layer.conv1.bias, relu=True, split=split)   # conv -> +bias -> ReLU
        self.conv2 = pk.pack_conv(layer.conv2.weight, layer.conv2.bias, split=split)
        f = lambda t: t.detach().float().contiguous()
        self.n1 = (f(layer.norm1.weight), f(layer.norm1.bias))
        self.n2 = (f(layer.norm2.weight), f(layer.norm2.bias))


class TransformerPack:
    """Packed weights of one ForwardTransformer (common_layers.py:188-223)."""

    def __init__(self, m, precise: bool = False):
        """precise: the split-precision form (every GEMM operand as fp16 [hi | lo | hi] column blocks, weights as
        [W_hi | W_hi | W_lo]) -- the duration predictor, whose output is rounded to integers (fast_pitch.py:174-177)."""
        self.d_model, self.heads = m.d_model, m.heads
        self.precise = bool(precise and eng.PRECISE_DUR and m.d_model % 8 == 0
                            and m.layers[0].conv1.weight.shape[0] % 8 == 0)
        self.layers = [FftLayerPack(l, split=self.precise) for l in m.layers]
        self.norm = (m.norm.weight.detach().float().contiguous(), m.norm.bias.detach().float().contiguous())
        self.pe = m.pos_encoder.pe.detach().float()[:, 0, :].contiguous()           # [max_len, E]
        self.scale = m.pos_encoder.scale.detach().float().contiguous()
        self.max_len = self.pe.shape[0]


def transformer_operand(tp: TransformerPack):
    """(column multiplier, dtype, round mode) of the GEMM-operand buffers of one ForwardTransformer: the split
    [hi | lo | hi] fp16 form for a precise (duration) transformer, else the plain fp16 operand copy (kind::f16 MMAs: twice
    the tensor rate and half the operand bytes of tf32 for the same 11-bit significand; LayerNorm outputs, attention
    outputs and ReLU'd conv1 outputs sit well inside fp16's range), else fp32 rounded to tf32 (FTB_F16_OPERANDS=0 or widths
    that are not multiples of 8).  The residual stream keeps its fp32 twins either way."""
    if tp.precise:
        return 3, torch.float16, 3
    if eng.F16_OPERANDS and tp.d_model % 8 == 0 and tp.layers[0].conv1.n % 8 == 0:
        return 1, torch.float16, 2
    return 1, torch.float32, True


def transformer_levels(prog: eng.Program, tag: str, tp: TransformerPack, x: torch.Tensor, out: torch.Tensor,
                       key_tok: Optional[torch.Tensor], round_final: bool = True,
                       x_full: Optional[torch.Tensor] = None):
    """Per-level closures of one ForwardTransformer over x [rows, E] (positional encoding already applied);
    the last level writes LayerNorm_final(LayerNorm2(...)) into `out`.  Post-norm FFT block
    (common_layers.py:170-185): x = LN1(x + MHA(x)); x = LN2(x + conv2(relu(conv1(x))))."""
    h, g = prog.h, prog.geo
    E = tp.d_model
    F = tp.layers[0].conv1.n
    PR = tp.precise                              # GEMM operands are [hi | lo | hi] fp16 buffers of 3x the width
    opw, opdt, rmode = transformer_operand(tp)
    assert x.dtype == opdt and x.shape[1] == opw * E and (opdt == torch.float32 or x_full is not None), (tag, x.dtype, x.shape)
    # head widths 64 / 128 (FastPitch): attention on the tensor cores (ftb_mha_split) over the in_proj output written as a
    # split [hi | lo | hi] fp16 operand; other widths (MultiFastPitch: 192 / 196 / 256) keep the fp32 SIMT kernel
    tc_attn = eng.TC_ATTENTION and (E // tp.heads) in (64, 128) and E % tp.heads == 0
    qkv = prog.buf(f'{tag}.qkv16', 9 * E, torch.float16) if tc_attn else prog.buf(f'{tag}.qkv', 3 * E)
    att = prog.buf(f'{tag}.att', opw * E, opdt)
    y = prog.buf(f'{tag}.blk_y', E)
    x1 = prog.buf(f'{tag}.x1', opw * E, opdt)
    x1f = prog.buf(f'{tag}.x1_full', E)          # unrounded twins: residual adds use full fp32 values
    c1 = prog.buf(f'{tag}.c1', opw * F, opdt)
    xs = [x, prog.buf(f'{tag}.x2', opw * E, opdt)]
    xsf = [x_full if x_full is not None else x, prog.buf(f'{tag}.x2_full', E)]
    levels = []

    def named(name, fn):
        def run():
            prog.label(f'{tag}.{name}')
            fn()
            prog.label('')
        return run

    n = len(tp.layers)
    for i, l in enumerate(tp.layers):
        src = xs[0] if i == 0 else xs[1]
        srcf = xsf[0] if i == 0 else xsf[1]
        dst = out if i == n - 1 else xs[1]
        dstf = None if i == n - 1 else xsf[1]
        last = tp.norm if i == n - 1 else (None, None)
        levels.append(named(f'l{i}.in_proj', lambda l=l, src=src: prog.gemm(src, l.in_proj, qkv, round_out=False,
                                                                            split_in=PR, split_out=tc_attn)))
        levels.append(lambda l=l: prog.call(
            lambda: (h.mha_split if tc_attn else h.mha)(qkv, E, tp.heads, key_tok, g.batch, g.seq_len, g.stride, att,
                                                        round_tf32=rmode),
            label=f'{tag}.mha', nbytes=4.0 * g.batch * g.seq_len * 4 * E))
        levels.append(named(f'l{i}.out_proj', lambda l=l, srcf=srcf: prog.gemm(att, l.out_proj, y, residual=srcf,
                                                                               round_out=False, split_in=PR)))
        levels.append(lambda l=l: prog.call(
            lambda: h.add_layernorm(y, None, E, E, l.n1[0], l.n1[1], None, None, g.batch, g.seq_len, g.stride, x1,
                                    round_tf32=rmode, out_full=x1f),
            label=f'{tag}.norm1', nbytes=8.0 * g.batch * g.seq_len * E))
        levels.append(named(f'l{i}.conv1', lambda l=l: prog.gemm(x1, l.conv1, c1, split_in=PR, split_out=PR)))
        levels.append(named(f'l{i}.conv2', lambda l=l: prog.gemm(c1, l.conv2, y, residual=x1f, round_out=False,
                                                                 split_in=PR)))
        # the last layer's output (`out`) is fp32: tf32-rounded when it feeds a GEMM, as computed for a predictor head
        levels.append(lambda l=l, dst=dst, dstf=dstf, last=last: prog.call(
            lambda: h.add_layernorm(y, None, E, E, l.n2[0], l.n2[1], last[0], last[1], g.batch, g.seq_len,
                                    g.stride, dst, round_tf32=((rmode if dst is not out else round_final)),
                                    out_full=dstf),
            label=f'{tag}.norm2', nbytes=8.0 * g.batch * g.seq_len * E))
    return levels
