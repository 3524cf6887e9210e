"""ForwardTacotron with the single-pass inference path on sm_100a kernels.

Drop-in for the reference's `models/forward_tacotron.py`: same constructor kwargs
(forward_tacotron.py:44-69), same state_dict (322 entries), same `generate` / `forward` /
`from_config` / `from_checkpoint` / `get_step` surface.  `generate()` and `_generate_mel()`
run entirely on libftb200.so (there is no PyTorch fallback on that path); the training-mode
`forward(batch)` stays stock PyTorch over the very same parameters.
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


class SeriesPredictor(nn.Module):
    """embedding -> 3x BatchNormConv(k=5) -> BiGRU -> Linear(1) -> /alpha."""

    def __init__(self, num_chars, emb_dim=64, conv_dims=256, rnn_dims=64, dropout=0.5):
        super().__init__()
        self.embedding = nn.Embedding(num_chars, emb_dim)
        dims = [emb_dim, conv_dims, conv_dims, conv_dims]
        self.convs = nn.ModuleList([BatchNormConv(dims[i], dims[i + 1], 5, relu=True) for i in range(3)])
        self.rnn = nn.GRU(conv_dims, rnn_dims, batch_first=True, bidirectional=True)
        self.lin = nn.Linear(2 * rnn_dims, 1)
        self.dropout = dropout

    def forward(self, x: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
        y = self.embedding(x).transpose(1, 2)
        for conv in self.convs:
            y = F.dropout(conv(y), p=self.dropout, training=self.training)
        y = self.rnn(y.transpose(1, 2))[0]
        return self.lin(y) / alpha


class ForwardTacotron(nn.Module, _accel.AcceleratedMixin):

    def __init__(self,
                 embed_dims: int, series_embed_dims: int, num_chars: int,
                 durpred_conv_dims: int, durpred_rnn_dims: int, durpred_dropout: float,
                 pitch_conv_dims: int, pitch_rnn_dims: int, pitch_dropout: float, pitch_strength: float,
                 energy_conv_dims: int, energy_rnn_dims: int, energy_dropout: float, energy_strength: float,
                 rnn_dims: int, prenet_dims: int, prenet_k: int, postnet_num_highways: int,
                 prenet_dropout: float, postnet_dims: int, postnet_k: int, prenet_num_highways: int,
                 postnet_dropout: float, n_mels: int, padding_value=-11.5129):
        ctor_kwargs = {k: v for k, v in locals().items() if k not in ('self', '__class__')}
        super().__init__()
        self._ctor_kwargs = ctor_kwargs            # forwardtacotron_b200.jit.script() rebuilds the module from these
        self.rnn_dims = rnn_dims
        self.padding_value = padding_value
        self.n_mels = n_mels
        self.embedding = nn.Embedding(num_chars, embed_dims)
        self.lr = LengthRegulator()
        self.dur_pred = SeriesPredictor(num_chars, series_embed_dims, durpred_conv_dims, durpred_rnn_dims,
                                        durpred_dropout)
        self.pitch_pred = SeriesPredictor(num_chars, series_embed_dims, pitch_conv_dims, pitch_rnn_dims,
                                          pitch_dropout)
        self.energy_pred = SeriesPredictor(num_chars, series_embed_dims, energy_conv_dims, energy_rnn_dims,
                                           energy_dropout)
        self.prenet = CBHG(K=prenet_k, in_channels=embed_dims, channels=prenet_dims,
                           proj_channels=[prenet_dims, embed_dims], num_highways=prenet_num_highways,
                           dropout=prenet_dropout)
        self.lstm = nn.LSTM(2 * prenet_dims, rnn_dims, batch_first=True, bidirectional=True)
        self.lin = nn.Linear(2 * rnn_dims, n_mels)
        self.register_buffer('step', torch.zeros(1, dtype=torch.long))
        self.postnet = CBHG(K=postnet_k, in_channels=n_mels, channels=postnet_dims,
                            proj_channels=[postnet_dims, n_mels], num_highways=postnet_num_highways,
                            dropout=postnet_dropout)
        self.post_proj = nn.Linear(2 * postnet_dims, n_mels, bias=False)
        self.pitch_strength = pitch_strength
        self.energy_strength = energy_strength
        self.pitch_proj = nn.Conv1d(1, 2 * prenet_dims, kernel_size=3, padding=1)
        self.energy_proj = nn.Conv1d(1, 2 * prenet_dims, kernel_size=3, padding=1)

    def __repr__(self):
        return f'ForwardTacotron, num params: {sum(int(np.prod(p.size())) for p in self.parameters())}'

    def get_step(self) -> int:
        return self.step.data.item()

    # ------------------------------------------------------------------ training path (stock torch)
    def forward(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        x, mel_len = batch['x'], batch['mel_len']
        target_len = batch['mel'].size(2)
        if self.training:
            self.step += 1
        if not self.training and x.is_cuda and not torch.is_grad_enabled():
            return self._forward_eval(batch)               # the GTA-dump / evaluation use (train_forward.py:33-51)

        dur_hat = self.dur_pred(x).squeeze(-1)
        pitch_hat = self.pitch_pred(x).transpose(1, 2)
        energy_hat = self.energy_pred(x).transpose(1, 2)

        h = self.prenet(self.embedding(x).transpose(1, 2))
        h = h + self.pitch_proj(batch['pitch'].unsqueeze(1)).transpose(1, 2) * self.pitch_strength
        h = h + self.energy_proj(batch['energy'].unsqueeze(1)).transpose(1, 2) * self.energy_strength
        h = self.lr(h, batch['dur'])
        packed = pack_padded_sequence(h, lengths=mel_len.cpu(), enforce_sorted=False, batch_first=True)
        h, _ = pad_packed_sequence(self.lstm(packed)[0], padding_value=self.padding_value, batch_first=True)
        mel = self.lin(h).transpose(1, 2)
        mel_post = self.post_proj(self.postnet(mel)).transpose(1, 2)
        return {'mel': self._pad(mel, target_len), 'mel_post': self._pad(mel_post, target_len),
                'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}

    def _pad(self, x: torch.Tensor, max_len: int) -> torch.Tensor:
        x = x[:, :, :max_len]
        return F.pad(x, [0, max_len - x.size(2), 0, 0], 'constant', self.padding_value)

    # ------------------------------------------------------------------ accelerated inference path
    def _build_packs(self) -> dict:
        return {
            'embedding': pk.round_tf32(self.embedding.weight),
            'embedding_full': self.embedding.weight.detach().float().contiguous(),
            'dur': _accel.pack_predictor(self.dur_pred, precise=True),
            'dur_emb': self.dur_pred.embedding.weight.detach().float().contiguous(),
            'pitch': _accel.pack_predictor(self.pitch_pred),
            'pitch_emb': pk.round_tf32(self.pitch_pred.embedding.weight),
            'energy': _accel.pack_predictor(self.energy_pred),
            'energy_emb': pk.round_tf32(self.energy_pred.embedding.weight),
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

    def _token_program(self, st: dict, batch: int, T: int, predictors: bool, prenet: bool) -> eng.Program:
        """Phoneme-side program: (optionally) the three SeriesPredictors and the embedding + CBHG prenet,
        level-interleaved so each persistent GEMM launch carries every branch's ready layer."""
        packs, h = st['packs'], st['handle']
        geo = eng.Geometry(batch, T)
        prog = eng.Program(h, geo)
        tok = prog.dense('tok', (batch, -1), torch.int64)
        branches, grus, heads = [], [], []
        if predictors:
            for name in ('dur', 'pitch', 'energy'):
                p = packs[name]
                emb = eng.predictor_emb_buffer(prog, name, p, packs[name + '_emb'].shape[1])
                prog.embed(tok, packs[name + '_emb'], emb, round_tf32=3 if p.precise else True)   # queued: one launch
                levels, gru, hbuf = eng.predictor_levels(prog, name, p, emb)
                branches.append(levels)
                grus.append(gru)
                heads.append((name, p, hbuf))
        if prenet:
            # the bank's operand copy of the embedding: fp16 (kind::f16 MMAs); the residual uses emb_full below
            E = packs['embedding'].shape[1]
            emb = prog.buf('emb', E, torch.float16 if (eng.F16_OPERANDS and E % 8 == 0) else torch.float32)
            prog.embed(tok, packs['embedding'], emb)
            embf = prog.buf('emb_full', emb.shape[1])         # unrounded copy: the CBHG residual (:114)
            prog.embed(tok, packs['embedding_full'], embf, round_tf32=False)
            out = prog.buf('prenet.h', 2 * packs['prenet'].rnn.hidden)
            levels, gru = eng.cbhg_levels(prog, 'prenet', packs['prenet'], emb, out, x_full=embf)
            branches.append(levels)
            grus.append(gru)
        eng.run_levels(prog, branches)
        prog.label('token_side_grus')
        prog.rnn(grus)
        prog.label('')
        for name, p, hbuf in heads:
            y = prog.dense(f'{name}.y', (batch, -1))
            prog.rowdot(hbuf, p.lin_w, p.lin_b, 1, (lambda: prog.alpha) if name == 'dur' else 1.0, y)
        prog.flush()                                                # the three heads: one launch
        prog.alpha = 1.0
        return prog

    def generate(self,
                 x: torch.Tensor,
                 alpha=1.0,
                 pitch_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x,
                 energy_function: Callable[[torch.Tensor], torch.Tensor] = lambda x: x) -> Dict[str, torch.Tensor]:
        self.eval()
        self._check_tokens(x)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('tok', B, T), lambda T: self._token_program(st, B, T, True, True))
            prog.bufs['tok'].copy_(x)
            prog.alpha = float(alpha)
            prog.run()
            dur_hat = prog.bufs['dur.y'].clone()                               # [B,T]  (dur_pred(x)/alpha).squeeze(2)
            pending = _accel.length_regulate_begin(st['handle'], dur_hat, apply_fallback=True)   # :176-177 (+ LR clamp)
            # the callbacks do not depend on the regulated length: they are enqueued while the scan result travels to the host
            pitch_hat = pitch_function(prog.bufs['pitch.y'].clone().unsqueeze(1))       # [B,1,T]
            energy_hat = energy_function(prog.bufs['energy.y'].clone().unsqueeze(1))
            pre = prog.bufs['prenet.h']
            if _accel.token_x_supported(st['packs'], pre.shape[1]):
                # the BiLSTM's x-projection per phoneme does not depend on the regulated length either
                _accel.token_xproj(st['handle'], st['packs'], prog, pre, None, 0, pitch_hat, energy_hat,
                                   float(self.pitch_strength), float(self.energy_strength))
            scan = pending.finish()
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, scan)

    def generate_jit(self, x: torch.Tensor, alpha: float = 1.0, beta: float = 1.0) -> Dict[str, torch.Tensor]:
        """The callback-free twin of generate() the reference exports to TorchScript (forward_tacotron.py:186-200):
        durations / alpha, pitch * beta, no pitch / energy callbacks and -- unlike generate() -- no eval() side effect.
        It runs on the same sm_100a kernels; the module itself is not scriptable (the kernels are reached through
        ctypes), so the method is offered under the reference's name for callers of an exported model."""
        if self.training:
            raise RuntimeError('generate_jit(): the module is in training mode; the reference exports and calls the '
                               'scripted module in eval mode (README "Export Model with TorchScript") -- call .eval() first')
        beta = float(beta)
        return self.generate(x, alpha=float(alpha), pitch_function=lambda p: p * beta)

    def _forward_eval(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """forward(batch) in eval mode under no_grad on the sm_100a kernels (forward_tacotron.py:118-165):
        predictors for the returned dur/pitch/energy, teacher-forced durations / pitch / energy for the mel branch,
        packed-sequence BiLSTM over batch['mel_len'] (tail = padding_value), outputs padded to the target length."""
        x = batch['x']
        self._check_tokens(x)
        st = self._accel_state(x.device)
        B, T = x.shape
        tok = self._program(st, ('tok', B, T), lambda T: self._token_program(st, B, T, True, True))
        tok.bufs['tok'].copy_(x)
        tok.alpha = 1.0
        tok.run()
        dur_hat = tok.bufs['dur.y'].clone()
        pitch_hat = tok.bufs['pitch.y'].clone().unsqueeze(1)
        energy_hat = tok.bufs['energy.y'].clone().unsqueeze(1)
        dur = batch['dur'].to(device=x.device, dtype=torch.float32).contiguous().clone()
        cum, lens, M_lr, _ = _accel.length_regulate(st['handle'], dur, apply_fallback=False)
        mel_len = batch['mel_len'].to(device=x.device, dtype=torch.int32).contiguous()
        M = int(mel_len.max())
        if M < 1 or M > M_lr or int(mel_len.min()) < 1:
            raise RuntimeError(f'forward(): mel_len must lie in [1, {M_lr}] (the regulated length), as '
                               'pack_padded_sequence requires (forward_tacotron.py:147)')
        pre = tok.bufs['prenet.h']
        prog = _accel.frame_program(self, st, 'frame_packed', tok, B, M, pre.shape[1], self.n_mels,
                                    packed_pad_value=float(self.padding_value))
        prog.bufs['mel_len'].copy_(mel_len)
        out = _accel.gather_and_run_frames(st['handle'], st['packs'], tok, prog, pre, None, 0, dur,
                                           batch['pitch'].to(x.device), batch['energy'].to(x.device), cum, lens, M,
                                           float(self.pitch_strength), float(self.energy_strength))
        target_len = batch['mel'].size(2)
        return {'mel': self._pad(out['mel'], target_len), 'mel_post': self._pad(out['mel_post'], target_len),
                'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}

    def _generate_mel(self, x: torch.Tensor, dur_hat: torch.Tensor, pitch_hat: torch.Tensor,
                      energy_hat: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Teacher-forced entry (forward_tacotron.py:205-234): durations/pitch/energy supplied by the caller."""
        self._check_tokens(x)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('prenet', B, T), lambda T: self._token_program(st, B, T, False, True))
            prog.bufs['tok'].copy_(x)
            prog.run()
            if not (dur_hat.is_cuda and dur_hat.dtype == torch.float32 and dur_hat.is_contiguous()):
                raise RuntimeError('_generate_mel(): dur_hat must be a contiguous fp32 CUDA tensor [B,T]')
            scan = _accel.length_regulate(st['handle'], dur_hat, apply_fallback=False)
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, scan)

    def _mel_from_prenet(self, st, tok_prog, dur_hat, pitch_hat, energy_hat, scan) -> Dict[str, torch.Tensor]:
        packs, h = st['packs'], st['handle']
        cum, lens, M, total = scan
        B, T = dur_hat.shape
        if M <= 0:
            raise RuntimeError('generate(): all rounded durations are zero -- nothing to synthesise')
        pre = tok_prog.bufs['prenet.h']
        C = pre.shape[1]
        prog = _accel.frame_program(self, st, 'frame', tok_prog, B, M, C, self.n_mels)
        return _accel.gather_and_run_frames(h, packs, tok_prog, prog, pre, None, 0, dur_hat, pitch_hat, energy_hat,
                                            cum, lens, M, float(self.pitch_strength), float(self.energy_strength))

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> 'ForwardTacotron':
        kwargs = config['forward_tacotron']['model']
        kwargs['num_chars'] = NUM_CHARS
        kwargs['n_mels'] = config['dsp']['num_mels']
        return cls(**kwargs)

    @classmethod
    def from_checkpoint(cls, path: Union[Path, str]) -> 'ForwardTacotron':
        ckpt = torch.load(path, map_location=torch.device('cpu'))
        model = cls.from_config(ckpt['config'])
        model.load_state_dict(ckpt['model'])
        return model
