"""Shared inference machinery of the accelerated models: weight-pack cache, program cache,
the frame-side (post length-regulator) program common to ForwardTacotron and
MultiForwardTacotron, and the host-side result assembly."""
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from .. import engine as eng
from .. import pack as pk
from .. import runtime as rt


def pack_bn_conv(m) -> pk.PackedGemm:
    bn = m.bnorm
    return pk.pack_bn_conv(m.conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var, m.relu is True)


def pack_cbhg(m) -> eng.CbhgPack:
    bank = [pack_bn_conv(c) for c in m.conv1d_bank]
    hws = [pk.pack_highway(h.W1.weight, h.W1.bias, h.W2.weight, h.W2.bias) for h in m.highways]
    return eng.CbhgPack(bank, pack_bn_conv(m.conv_project1), pack_bn_conv(m.conv_project2),
                        pk.pack_linear(m.pre_highway.weight), hws, pk.PackedRnn(m.rnn),
                        m.conv1d_bank[0].conv.in_channels, m.conv1d_bank[0].conv.out_channels)


def pack_predictor(m) -> eng.PredictorPack:
    return eng.PredictorPack([pack_bn_conv(c) for c in m.convs], pk.PackedRnn(m.rnn), m.lin.weight, m.lin.bias)


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
            handle = rt.Handle.get(device)
            st = {'sig': sig, 'device': device, 'handle': handle, 'packs': self._build_packs(),
                  'programs': OrderedDict()}
            self.__dict__['_accel'] = st
        return st

    def _program(self, st: dict, key, builder):
        progs = st['programs']
        if key in progs:
            progs.move_to_end(key)
            return progs[key]
        prog = builder()
        progs[key] = prog
        while len(progs) > self.MAX_PROGRAMS * 2:
            _, old = progs.popitem(last=False)
            old.close()
        return prog

    @staticmethod
    def _check_tokens(x: torch.Tensor):
        if not x.is_cuda:
            raise RuntimeError('generate(): x must be a CUDA tensor -- this package has no CPU path '
                               '(the sm_100a kernels are the implementation)')
        if x.dim() != 2 or x.dtype != torch.int64:
            raise RuntimeError(f'generate(): x must be int64 [batch, tokens], got {x.dtype} {tuple(x.shape)}')


def build_frame_program(handle: rt.Handle, packs: dict, batch: int, M: int, in_channels: int, n_mels: int):
    """LR output [B,M,C] -> BiLSTM -> lin -> CBHG postnet -> post_proj
    (forward_tacotron.py:224-231).  Returns the Program; the LR gather writes prog.buf('lr')."""
    geo = eng.Geometry(batch, M)
    prog = eng.Program(handle, geo)
    lstm: pk.PackedRnn = packs['lstm']
    x = prog.buf('lr', in_channels)
    xp = prog.buf('lstm.xproj', lstm.ih.n)
    hl = prog.buf('lstm.h', 2 * lstm.hidden)
    mel_cl = prog.buf('mel_cl', n_mels)
    mel = prog.dense('mel', (batch, n_mels, M))
    mel_post = prog.dense('mel_post', (batch, n_mels, M))
    post = prog.buf('postnet.h', 2 * packs['postnet'].rnn.hidden)

    prog.label('lstm_ih')
    prog.gemm(x, lstm.ih, xp, round_out=False)                    # W_ih half of self.lstm (:224)
    prog.label('lstm')
    prog.rnn([(lstm, xp, hl, 0)])
    prog.label('lin')
    prog.gemm(hl, packs['lin'], mel_cl, out_t=mel)                # self.lin + transpose (:226-227)
    prog.flush()
    prog.label('')
    levels, gru = eng.cbhg_levels(prog, 'postnet', packs['postnet'], mel_cl, post)
    eng.run_levels(prog, [levels])
    prog.label('postnet.gru')
    prog.rnn([gru])
    prog.label('post_proj')
    prog.gemm(post, packs['post_proj'], None, out_t=mel_post)     # self.post_proj + transpose (:230-231)
    prog.flush()
    return prog


def length_regulate(handle: rt.Handle, dur: torch.Tensor, apply_fallback: bool):
    """Runs K5's scan on `dur` (clamped / fallback-filled IN PLACE like the reference) and reads the
    three integers the host needs (max length, fallback flag, total frames): the one D2H sync."""
    B, T = dur.shape
    dev = dur.device
    cum = torch.empty((B, T), dtype=torch.int32, device=dev)
    lens = torch.empty((B,), dtype=torch.int32, device=dev)
    meta = torch.empty((4,), dtype=torch.int32, device=dev)
    handle.lr_scan(dur, B, T, apply_fallback, cum, lens, meta)
    m = meta.cpu()
    return cum, lens, int(m[0]), int(m[2])
