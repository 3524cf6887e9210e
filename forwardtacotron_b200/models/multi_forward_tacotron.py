"""MultiForwardTacotron (speaker-conditioned) with the single-pass inference path on sm_100a kernels.

Drop-in for the reference's `models/multi_forward_tacotron.py`: same constructor kwargs
(multi_forward_tacotron.py:98-129), same state_dict, same `generate(x, speaker_emb, alpha,
pitch_function, energy_function)` surface (:243-248).  `generate()` runs on libftb200.so only.

Batch semantics (documented deviation): the reference computes the pitch
condition as `argmax(pitch_cond_hat.squeeze(), dim=1).unsqueeze(0)` (:252), which is well-formed for
batch 1 only (IndexError at batch >= 2).  Here the arg-max is taken over the class axis per (b, t);
at batch 1 this is the same tensor, and the returned shapes match the reference's test
(tests/test_multi_forward_tacotron.py:42-49): 'pitch_cond' is [B, 1, T].
"""
from pathlib import Path
from typing import Any, Callable, Dict, Union

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence

from .. import engine as eng
from .. import pack as pk
from .. import runtime as rt
from ..configs import NUM_CHARS
from . import _accel
from .common_layers import CBHG, BatchNormConv, LengthRegulator


def _with_speaker(e: torch.Tensor, semb: torch.Tensor) -> torch.Tensor:
    return torch.cat([e, semb[:, None, :].expand(-1, e.size(1), -1)], dim=2)


class SeriesPredictor(nn.Module):
    """[embedding | speaker] -> 3x BatchNormConv(k=5) -> BiGRU -> Linear(out_dim) -> /alpha."""

    def __init__(self, num_chars: int, emb_dim: int = 64, conv_dims: int = 256, rnn_dims: int = 64,
                 dropout: float = 0.5, speaker_emb_dims: int = 256, out_dim: int = 1):
        super().__init__()
        self.embedding = nn.Embedding(num_chars, emb_dim)
        dims = [emb_dim + speaker_emb_dims, conv_dims, conv_dims, conv_dims]
        self.convs = nn.ModuleList([BatchNormConv(dims[i], dims[i + 1], 5, relu=True) for i in range(3)])
        self.rnn = nn.GRU(conv_dims, rnn_dims, batch_first=True, bidirectional=True)
        self.lin = nn.Linear(2 * rnn_dims, out_dim)
        self.dropout = dropout

    def _tail(self, y: torch.Tensor, alpha: float) -> torch.Tensor:
        y = y.transpose(1, 2)
        for conv in self.convs:
            y = F.dropout(conv(y), p=self.dropout, training=self.training)
        y = self.rnn(y.transpose(1, 2))[0]
        return self.lin(y) / alpha

    def forward(self, x: torch.Tensor, semb: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
        return self._tail(_with_speaker(self.embedding(x), semb), alpha)


class ConditionalSeriesPredictor(SeriesPredictor):
    """[embedding | pitch-condition embedding | speaker] -> same tail."""

    def __init__(self, num_chars: int, emb_dim: int = 64, cond_emb_size: int = 4, cond_emb_dims: int = 8,
                 conv_dims: int = 256, rnn_dims: int = 64, dropout: float = 0.5, speaker_emb_dims: int = 256):
        nn.Module.__init__(self)
        self.embedding = nn.Embedding(num_chars, emb_dim)
        self.pitch_cond_embedding = nn.Embedding(cond_emb_size, cond_emb_dims)
        dims = [emb_dim + cond_emb_dims + speaker_emb_dims, conv_dims, conv_dims, conv_dims]
        self.convs = nn.ModuleList([BatchNormConv(dims[i], dims[i + 1], 5, relu=True) for i in range(3)])
        self.rnn = nn.GRU(conv_dims, rnn_dims, batch_first=True, bidirectional=True)
        self.lin = nn.Linear(2 * rnn_dims, 1)
        self.dropout = dropout

    def forward(self, x: torch.Tensor, x_cond: torch.Tensor, speaker_emb: torch.Tensor,
                alpha: float = 1.0) -> torch.Tensor:
        e = torch.cat([self.embedding(x), self.pitch_cond_embedding(x_cond)], dim=2)
        return self._tail(_with_speaker(e, speaker_emb), alpha)


class MultiForwardTacotron(nn.Module, _accel.AcceleratedMixin):

    def __init__(self,
                 embed_dims: int, series_embed_dims: int, num_chars: int,
                 durpred_conv_dims: int, durpred_rnn_dims: int, durpred_dropout: float,
                 pitch_conv_dims: int, pitch_rnn_dims: int, pitch_dropout: float, pitch_strength: float,
                 pitch_cond_conv_dims: int, pitch_cond_rnn_dims: int, pitch_cond_dropout: float,
                 energy_conv_dims: int, energy_rnn_dims: int, energy_dropout: float, energy_strength: float,
                 rnn_dims: int, prenet_dims: int, prenet_k: int, postnet_num_highways: int,
                 prenet_dropout: float, postnet_dims: int, postnet_k: int, prenet_num_highways: int,
                 postnet_dropout: float, n_mels: int, speaker_emb_dims: int, pitch_cond_emb_dims: int,
                 pitch_cond_categorical_dims: int, padding_value=-11.5129):
        super().__init__()
        self.rnn_dims = rnn_dims
        self.padding_value = padding_value
        self.n_mels = n_mels
        self.speaker_emb_dims = speaker_emb_dims
        self.embedding = nn.Embedding(num_chars, embed_dims)
        self.lr = LengthRegulator()
        self.dur_pred = ConditionalSeriesPredictor(num_chars=num_chars, emb_dim=series_embed_dims,
                                                   conv_dims=durpred_conv_dims, rnn_dims=durpred_rnn_dims,
                                                   cond_emb_dims=pitch_cond_emb_dims, dropout=durpred_dropout)
        self.pitch_cond_pred = SeriesPredictor(num_chars=num_chars, emb_dim=series_embed_dims,
                                               conv_dims=pitch_cond_conv_dims, rnn_dims=pitch_cond_rnn_dims,
                                               dropout=pitch_cond_dropout, out_dim=pitch_cond_categorical_dims)
        self.pitch_pred = ConditionalSeriesPredictor(num_chars=num_chars, emb_dim=series_embed_dims,
                                                     conv_dims=pitch_conv_dims, rnn_dims=pitch_rnn_dims,
                                                     cond_emb_dims=pitch_cond_emb_dims, dropout=pitch_dropout)
        self.energy_pred = SeriesPredictor(num_chars=num_chars, emb_dim=series_embed_dims,
                                           conv_dims=energy_conv_dims, rnn_dims=energy_rnn_dims,
                                           dropout=energy_dropout)
        self.prenet = CBHG(K=prenet_k, in_channels=embed_dims, channels=prenet_dims,
                           proj_channels=[prenet_dims, embed_dims], num_highways=prenet_num_highways,
                           dropout=prenet_dropout)
        self.lstm = nn.LSTM(2 * prenet_dims + speaker_emb_dims, rnn_dims, batch_first=True, bidirectional=True)
        self.lin = nn.Linear(2 * rnn_dims, n_mels)
        self.register_buffer('step', torch.zeros(1, dtype=torch.long))
        self.postnet = CBHG(K=postnet_k, in_channels=n_mels, channels=postnet_dims,
                            proj_channels=[postnet_dims, n_mels], num_highways=postnet_num_highways,
                            dropout=postnet_dropout)
        self.post_proj = nn.Linear(2 * postnet_dims, n_mels, bias=False)
        self.pitch_strength = pitch_strength
        self.energy_strength = energy_strength
        self.pitch_proj = nn.Conv1d(1, 2 * prenet_dims + speaker_emb_dims, kernel_size=3, padding=1)
        self.energy_proj = nn.Conv1d(1, 2 * prenet_dims + speaker_emb_dims, kernel_size=3, padding=1)

    def __repr__(self):
        return f'MultiForwardTacotron, num params: {sum(int(np.prod(p.size())) for p in self.parameters())}'

    def get_step(self) -> int:
        return self.step.data.item()

    # ------------------------------------------------------------------ training path (stock torch)
    def forward(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        x, semb, mel_len = batch['x'], batch['speaker_emb'], batch['mel_len']
        target_len = batch['mel'].size(2)
        if self.training:
            self.step += 1
        if not self.training and x.is_cuda and not torch.is_grad_enabled():
            return self._forward_eval(batch)               # the GTA-dump / evaluation use (train_forward.py:33-51)
        pitch_cond_hat = self.pitch_cond_pred(x, semb).squeeze(-1)
        dur_hat = self.dur_pred(x, batch['pitch_cond'], semb).squeeze(-1)
        pitch_hat = self.pitch_pred(x, batch['pitch_cond'], semb).transpose(1, 2)
        energy_hat = self.energy_pred(x, semb).transpose(1, 2)

        h = _with_speaker(self.prenet(self.embedding(x).transpose(1, 2)), semb)
        h = h + self.pitch_proj(batch['pitch'].unsqueeze(1)).transpose(1, 2) * self.pitch_strength
        h = h + self.energy_proj(batch['energy'].unsqueeze(1)).transpose(1, 2) * self.energy_strength
        h = self.lr(h, batch['dur'])
        packed = pack_padded_sequence(h, lengths=mel_len.cpu(), enforce_sorted=False, batch_first=True)
        h, _ = pad_packed_sequence(self.lstm(packed)[0], padding_value=self.padding_value, batch_first=True)
        mel = self.lin(h).transpose(1, 2)
        mel_post = self.post_proj(self.postnet(mel)).transpose(1, 2)
        return {'mel': self._pad(mel, target_len), 'mel_post': self._pad(mel_post, target_len),
                'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat, 'pitch_cond': pitch_cond_hat}

    def _pad(self, x: torch.Tensor, max_len: int) -> torch.Tensor:
        x = x[:, :, :max_len]
        return F.pad(x, [0, max_len - x.size(2), 0, 0], 'constant', self.padding_value)

    # ------------------------------------------------------------------ accelerated inference path
    def _build_packs(self) -> dict:
        packs = {
            'embedding': pk.round_tf32(self.embedding.weight),
            'embedding_full': self.embedding.weight.detach().float().contiguous(),
            'prenet': _accel.pack_cbhg(self.prenet),
            'lstm': pk.PackedRnn(self.lstm),
            'lin': pk.pack_linear(self.lin.weight, self.lin.bias),
            # lin(padding vector) in exact fp32: the mel rows behind a packed sentence (eval-mode forward only)
            'lin_pad_vec': (self.lin.weight.detach().float().sum(dim=1) * float(self.padding_value)
                            + self.lin.bias.detach().float()).contiguous(),
            'postnet': _accel.pack_cbhg(self.postnet, split_bank=True),
            'post_proj': pk.pack_linear(self.post_proj.weight),
            'pitch_proj': (self.pitch_proj.weight.detach().float().reshape(-1, 3).contiguous(),
                           self.pitch_proj.bias.detach().float().contiguous()),
            'energy_proj': (self.energy_proj.weight.detach().float().reshape(-1, 3).contiguous(),
                            self.energy_proj.bias.detach().float().contiguous()),
        }
        for name, mod in (('dur', self.dur_pred), ('pitch', self.pitch_pred), ('energy', self.energy_pred),
                          ('pitch_cond', self.pitch_cond_pred)):
            # the two predictors whose output is rounded to integers (durations :253-255 / LengthRegulator, and the
            # arg-max of the pitch condition :251-252) run the split-precision path; their embedding kernel rounds
            packs[name] = _accel.pack_predictor(mod, precise=name in ('dur', 'pitch_cond'))
            packs[name + '_emb'] = mod.embedding.weight.detach().float().contiguous()
            if hasattr(mod, 'pitch_cond_embedding'):
                packs[name + '_cemb'] = mod.pitch_cond_embedding.weight.detach().float().contiguous()
        return packs

    def _token_program(self, st: dict, batch: int, T: int, predictors: bool, teacher_cond: bool = False) -> eng.Program:
        """Phoneme-side program.  Phase 1: pitch_cond_pred, energy_pred and the CBHG prenet (level
        interleaved) + their GRUs; arg-max of the pitch condition (:251-252).  Phase 2: dur_pred and
        pitch_pred, which consume the pitch condition (:253,256).  `teacher_cond` (eval-mode forward, :199-200):
        the condition is the caller's batch['pitch_cond'] -- the 'pitch_cond' buffer is filled before run() and the
        arg-max is left out."""
        packs, h = st['packs'], st['handle']
        geo = eng.Geometry(batch, T)
        prog = eng.Program(h, geo)
        S = self.speaker_emb_dims
        tok = prog.dense('tok', (batch, -1), torch.int64)
        semb = prog.dense('semb', (batch, S))
        pcond = prog.dense('pitch_cond', (batch, -1), torch.int64)
        heads = []

        def add_predictor(name, conditional):
            p = packs[name]
            e1 = packs[name + '_emb'].shape[1]
            e2 = packs[name + '_cemb'].shape[1] if conditional else 0
            width = e1 + e2 + S
            emb = eng.predictor_emb_buffer(prog, name, p, pk.ceil_to(width, 4))
            prog.embed(tok, packs[name + '_emb'], emb, tok2=pcond if conditional else None,
                       table2=packs[name + '_cemb'] if conditional else None, extra=semb,
                       round_tf32=3 if p.precise else True)
            levels, gru, hbuf = eng.predictor_levels(prog, name, p, emb)
            heads.append((name, p, hbuf))
            return levels, gru

        branches, grus = [], []
        if predictors:
            for name in ('pitch_cond', 'energy'):
                levels, gru = add_predictor(name, False)
                branches.append(levels)
                grus.append(gru)
        # the bank's operand copy of the embedding: fp16 (kind::f16 MMAs); the residual uses emb_full below
        E = packs['embedding'].shape[1]
        emb = prog.buf('emb', E, torch.float16 if (eng.F16_OPERANDS and E % 8 == 0) else torch.float32)
        prog.embed(tok, packs['embedding'], emb)
        embf = prog.buf('emb_full', emb.shape[1])             # unrounded copy: the CBHG residual
        prog.embed(tok, packs['embedding_full'], embf, round_tf32=False)
        out = prog.buf('prenet.h', 2 * packs['prenet'].rnn.hidden)
        levels, gru = eng.cbhg_levels(prog, 'prenet', packs['prenet'], emb, out, x_full=embf)
        branches.append(levels)
        grus.append(gru)
        eng.run_levels(prog, branches)
        prog.label('token_side_grus')
        prog.rnn(grus)
        prog.label('')

        def add_head(name, p, hbuf):
            n_out = p.lin_w.shape[0]
            y = prog.dense(f'{name}.y', (batch, -1, n_out))
            prog.call(lambda: h.rowdot(hbuf, hbuf.shape[1], hbuf.shape[1], p.lin_w, p.lin_b, n_out,
                                       prog.alpha if name == 'dur' else 1.0, batch, geo.seq_len, geo.stride, y),
                      label=f'{name}.lin')
            return y

        if predictors:
            done = list(heads)
            heads.clear()
            logits = None
            for name, p, hbuf in done:
                y = add_head(name, p, hbuf)
                if name == 'pitch_cond':
                    logits = y
            n_cls = logits.shape[2]
            if not teacher_cond:
                prog.call(lambda: h.argmax(logits, batch * geo.seq_len, n_cls, pcond), label='pitch_cond.argmax')
            branches2, grus2 = [], []
            for name in ('dur', 'pitch'):
                levels, gru = add_predictor(name, True)
                branches2.append(levels)
                grus2.append(gru)
            eng.run_levels(prog, branches2)
            prog.label('conditional_grus')
            prog.rnn(grus2)
            prog.label('')
            for name, p, hbuf in heads:
                add_head(name, p, hbuf)
        prog.alpha = 1.0
        return prog

    @staticmethod
    def _check_speaker(x: torch.Tensor, speaker_emb: torch.Tensor, dims: int):
        if not (speaker_emb.is_cuda and speaker_emb.dim() == 2 and speaker_emb.shape == (x.shape[0], dims)):
            raise RuntimeError(f'generate(): speaker_emb must be a CUDA tensor [batch, {dims}], got '
                               f'{tuple(speaker_emb.shape)} on {speaker_emb.device}')

    def generate(self,
                 x: torch.Tensor,
                 speaker_emb: torch.Tensor,
                 alpha=1.0,
                 pitch_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x,
                 energy_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x) -> Dict[str, torch.Tensor]:
        self.eval()
        self._check_tokens(x)
        self._check_speaker(x, speaker_emb, self.speaker_emb_dims)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('tok', B, T), lambda T: self._token_program(st, B, T, True))
            prog.bufs['tok'].copy_(x)
            prog.bufs['semb'].copy_(speaker_emb.to(torch.float32))
            prog.alpha = float(alpha)
            prog.run()
            pitch_cond_hat = prog.bufs['pitch_cond'].clone()                   # [B,T] int64
            dur_hat = prog.bufs['dur.y'].reshape(B, T).clone()
            pending = _accel.length_regulate_begin(st['handle'], dur_hat, apply_fallback=True)   # :254-255 (+ LR clamp)
            # the callbacks do not depend on the regulated length: they are enqueued while the scan result travels to the host
            pitch_hat = pitch_function(prog.bufs['pitch.y'].reshape(B, 1, T).clone())
            energy_hat = energy_function(prog.bufs['energy.y'].reshape(B, 1, T).clone())
            pre, C2 = prog.bufs['prenet.h'], self.speaker_emb_dims
            if _accel.token_x_supported(st['packs'], pre.shape[1] + C2):
                # the BiLSTM's x-projection per phoneme does not depend on the regulated length either
                _accel.token_xproj(st['handle'], st['packs'], prog, pre, prog.bufs['semb'], C2, pitch_hat, energy_hat,
                                   float(self.pitch_strength), float(self.energy_strength))
            scan = pending.finish()
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, pitch_cond_hat, scan)

    def _forward_eval(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """forward(batch) in eval mode under no_grad on the sm_100a kernels (multi_forward_tacotron.py:186-241):
        predictors for the returned dur / pitch / energy / pitch_cond (dur and pitch conditioned on the TARGET pitch
        condition, :199-200), teacher-forced mel branch with the packed-sequence BiLSTM over batch['mel_len']."""
        x, semb = batch['x'], batch['speaker_emb']
        self._check_tokens(x)
        self._check_speaker(x, semb, self.speaker_emb_dims)
        st = self._accel_state(x.device)
        B, T = x.shape
        tok = self._program(st, ('tok_teacher', B, T), lambda T: self._token_program(st, B, T, True, teacher_cond=True))
        tok.bufs['tok'].copy_(x)
        tok.bufs['semb'].copy_(semb.to(torch.float32))
        tok.bufs['pitch_cond'].copy_(batch['pitch_cond'])
        tok.alpha = 1.0
        tok.run()
        dur_hat = tok.bufs['dur.y'].reshape(B, T).clone()
        pitch_hat = tok.bufs['pitch.y'].reshape(B, 1, T).clone()
        energy_hat = tok.bufs['energy.y'].reshape(B, 1, T).clone()
        pitch_cond_hat = tok.bufs['pitch_cond.y'].clone()                      # [B,T,3] logits
        dur = batch['dur'].to(device=x.device, dtype=torch.float32).contiguous().clone()
        cum, lens, M_lr, _ = _accel.length_regulate(st['handle'], dur, apply_fallback=False)
        mel_len = batch['mel_len'].to(device=x.device, dtype=torch.int32).contiguous()
        M = int(mel_len.max())
        if M < 1 or M > M_lr or int(mel_len.min()) < 1:
            raise RuntimeError(f'forward(): mel_len must lie in [1, {M_lr}] (the regulated length), as '
                               'pack_padded_sequence requires (multi_forward_tacotron.py:222)')
        pre = tok.bufs['prenet.h']
        C1, C2 = pre.shape[1], self.speaker_emb_dims
        prog = _accel.frame_program(self, st, 'frame_packed', tok, B, M, C1 + C2, self.n_mels,
                                    packed_pad_value=float(self.padding_value))
        prog.bufs['mel_len'].copy_(mel_len)
        out = _accel.gather_and_run_frames(st['handle'], st['packs'], tok, prog, pre, tok.bufs['semb'], C2, dur,
                                           batch['pitch'].to(x.device), batch['energy'].to(x.device), cum, lens, M,
                                           float(self.pitch_strength), float(self.energy_strength))
        target_len = batch['mel'].size(2)
        return {'mel': self._pad(out['mel'], target_len), 'mel_post': self._pad(out['mel_post'], target_len),
                'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat, 'pitch_cond': pitch_cond_hat}

    def _generate_mel(self, x: torch.Tensor, semb: torch.Tensor, dur_hat: torch.Tensor, pitch_hat: torch.Tensor,
                      pitch_cond_hat: torch.Tensor, energy_hat: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Teacher-forced entry (multi_forward_tacotron.py:270-305)."""
        self._check_tokens(x)
        self._check_speaker(x, semb, self.speaker_emb_dims)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('prenet', B, T), lambda T: self._token_program(st, B, T, False))
            prog.bufs['tok'].copy_(x)
            prog.bufs['semb'].copy_(semb.to(torch.float32))
            prog.run()
            if not (dur_hat.is_cuda and dur_hat.dtype == torch.float32 and dur_hat.is_contiguous()):
                raise RuntimeError('_generate_mel(): dur_hat must be a contiguous fp32 CUDA tensor [B,T]')
            scan = _accel.length_regulate(st['handle'], dur_hat, apply_fallback=False)
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, pitch_cond_hat, scan)

    def _mel_from_prenet(self, st, tok_prog, dur_hat, pitch_hat, energy_hat, pitch_cond_hat, scan):
        packs, h = st['packs'], st['handle']
        cum, lens, M, total = scan
        B, T = dur_hat.shape
        if M <= 0:
            raise RuntimeError('generate(): all rounded durations are zero -- nothing to synthesise')
        pre = tok_prog.bufs['prenet.h']
        C1, C2 = pre.shape[1], self.speaker_emb_dims
        prog = _accel.frame_program(self, st, 'frame', tok_prog, B, M, C1 + C2, self.n_mels)
        out = _accel.gather_and_run_frames(h, packs, tok_prog, prog, pre, tok_prog.bufs['semb'], C2, dur_hat,
                                           pitch_hat, energy_hat, cum, lens, M,
                                           float(self.pitch_strength), float(self.energy_strength))
        out['pitch_cond'] = pitch_cond_hat.reshape(B, T).unsqueeze(1)          # :303-305
        return out

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> 'MultiForwardTacotron':
        kwargs = config['multi_forward_tacotron']['model']
        kwargs['num_chars'] = NUM_CHARS
        kwargs['n_mels'] = config['dsp']['num_mels']
        return cls(**kwargs)

    @classmethod
    def from_checkpoint(cls, path: Union[Path, str]) -> 'MultiForwardTacotron':
        ckpt = torch.load(path, map_location=torch.device('cpu'))
        model = cls.from_config(ckpt['config'])
        model.load_state_dict(ckpt['model'])
        return model
