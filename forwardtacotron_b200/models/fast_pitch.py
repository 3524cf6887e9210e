"""FastPitch (FFT-block variant) with the single-pass inference path on sm_100a kernels.

Drop-in for the reference's `models/fast_pitch.py`: same constructor kwargs (fast_pitch.py:46-77), same
state_dict, same `generate(x, alpha, pitch_function, energy_function)` surface (:167-171).  The
QKV / out_proj / conv1(k=9) / conv2(k=1) / mel layers run on the tcgen05 conv-GEMM kernel; attention,
residual+LayerNorm and the positional encoding are the K7/K8 kernels (csrc/ftb_fft.cu).

Reference semantics kept on purpose: the predictors run WITHOUT a key-padding mask in
generate() (:174-180); the prenet masks keys where the token id is 0 (:200); the postnet attends over all
M frames, padded ones included (:215); 'mel' and 'mel_post' are the same tensor (:220); sequences longer
than PositionalEncoding.max_len (5000) are an error, as in the reference (common_layers.py:129,144).
"""
from pathlib import Path
from typing import Any, Callable, Dict, Optional, Union

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import engine as eng
from .. import pack as pk
from ..configs import NUM_CHARS
from . import _accel
from .common_layers import ForwardTransformer, LengthRegulator, make_token_len_mask


class SeriesPredictor(nn.Module):
    """embedding -> ForwardTransformer -> Linear(1) -> /alpha."""

    def __init__(self, num_chars: int, d_model: int, n_heads: int, d_fft: int, layers: int,
                 conv1_kernel: int, conv2_kernel: int, dropout=0.1):
        super().__init__()
        self.embedding = nn.Embedding(num_chars, d_model)
        self.transformer = ForwardTransformer(heads=n_heads, dropout=dropout, d_model=d_model, d_fft=d_fft,
                                              conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel, layers=layers)
        self.lin = nn.Linear(d_model, 1)

    def forward(self, x: torch.Tensor, src_pad_mask: Optional[torch.Tensor] = None,
                alpha: float = 1.0) -> torch.Tensor:
        y = self.transformer(self.embedding(x), src_pad_mask=src_pad_mask)
        return self.lin(y) / alpha


class FastPitch(nn.Module, _accel.AcceleratedMixin):

    def __init__(self,
                 num_chars: int,
                 durpred_dropout: float, durpred_d_model: int, durpred_n_heads: int, durpred_layers: int,
                 durpred_d_fft: int,
                 pitch_dropout: float, pitch_d_model: int, pitch_n_heads: int, pitch_layers: int, pitch_d_fft: int,
                 energy_dropout: float, energy_d_model: int, energy_n_heads: int, energy_layers: int,
                 energy_d_fft: int,
                 pitch_strength: float, energy_strength: float,
                 d_model: int, conv1_kernel: int, conv2_kernel: int,
                 prenet_layers: int, prenet_heads: int, prenet_fft: int, prenet_dropout: float,
                 postnet_layers: int, postnet_heads: int, postnet_fft: int, postnet_dropout: float,
                 n_mels: int, padding_value=-11.5129):
        super().__init__()
        self.padding_value = padding_value
        self.n_mels = n_mels
        self.lr = LengthRegulator()
        kern = dict(conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel)
        self.dur_pred = SeriesPredictor(num_chars=num_chars, d_model=durpred_d_model, n_heads=durpred_n_heads,
                                        layers=durpred_layers, d_fft=durpred_d_fft, dropout=durpred_dropout, **kern)
        self.pitch_pred = SeriesPredictor(num_chars=num_chars, d_model=pitch_d_model, n_heads=pitch_n_heads,
                                          layers=pitch_layers, d_fft=pitch_d_fft, dropout=pitch_dropout, **kern)
        self.energy_pred = SeriesPredictor(num_chars=num_chars, d_model=energy_d_model, n_heads=energy_n_heads,
                                           layers=energy_layers, d_fft=energy_d_fft, dropout=energy_dropout, **kern)
        self.embedding = nn.Embedding(num_embeddings=num_chars, embedding_dim=d_model)
        self.prenet = ForwardTransformer(heads=prenet_heads, dropout=prenet_dropout, d_model=d_model,
                                         d_fft=prenet_fft, layers=prenet_layers, **kern)
        self.postnet = ForwardTransformer(heads=postnet_heads, dropout=postnet_dropout, d_model=d_model,
                                          d_fft=postnet_fft, layers=postnet_layers, **kern)
        self.lin = nn.Linear(d_model, n_mels)
        self.register_buffer('step', torch.zeros(1, dtype=torch.long))
        self.pitch_strength = pitch_strength
        self.energy_strength = energy_strength
        self.pitch_proj = nn.Conv1d(1, d_model, kernel_size=3, padding=1)
        self.energy_proj = nn.Conv1d(1, d_model, kernel_size=3, padding=1)

    def __repr__(self):
        return f'FastPitch, num params: {sum(int(np.prod(p.size())) for p in self.parameters())}'

    def get_step(self) -> int:
        return self.step.data.item()

    # ------------------------------------------------------------------ training path (stock torch)
    def forward(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        x, mel_lens = batch['x'], batch['mel_len']
        target_len = batch['mel'].size(2)
        if self.training:
            self.step += 1
        if not self.training and x.is_cuda and not torch.is_grad_enabled():
            return self._forward_eval(batch)               # the GTA-dump / evaluation use (train_forward.py:33-51)
        tok_mask = make_token_len_mask(x.transpose(0, 1))
        dur_hat = self.dur_pred(x, src_pad_mask=tok_mask).squeeze(-1)
        pitch_hat = self.pitch_pred(x, src_pad_mask=tok_mask).transpose(1, 2)
        energy_hat = self.energy_pred(x, src_pad_mask=tok_mask).transpose(1, 2)
        h = self.prenet(self.embedding(x), src_pad_mask=tok_mask)
        h = h + self.pitch_proj(batch['pitch'].unsqueeze(1)).transpose(1, 2) * self.pitch_strength
        h = h + self.energy_proj(batch['energy'].unsqueeze(1)).transpose(1, 2) * self.energy_strength
        h = self.lr(h, batch['dur'])
        frame_mask = torch.arange(h.size(1), device=h.device)[None, :] >= mel_lens.to(h.device)[:, None]
        h = self.postnet(h, src_pad_mask=frame_mask)
        mel = self.pad(self.lin(h).transpose(1, 2), target_len)
        return {'mel': mel, 'mel_post': mel.clone(), 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}

    def pad(self, x: torch.Tensor, max_len: int) -> torch.Tensor:
        x = x[:, :, :max_len]
        return F.pad(x, [0, max_len - x.size(2), 0, 0], 'constant', self.padding_value)

    # ------------------------------------------------------------------ accelerated inference path
    def _build_packs(self) -> dict:
        f = lambda t: t.detach().float().contiguous()   # noqa: E731
        packs = {
            'embedding': f(self.embedding.weight),
            'prenet': _accel.TransformerPack(self.prenet),
            'postnet': _accel.TransformerPack(self.postnet),
            'lin': pk.pack_linear(self.lin.weight, self.lin.bias),
            'pitch_proj': (f(self.pitch_proj.weight).reshape(-1, 3).contiguous(), f(self.pitch_proj.bias)),
            'energy_proj': (f(self.energy_proj.weight).reshape(-1, 3).contiguous(), f(self.energy_proj.bias)),
        }
        for name, mod in (('dur', self.dur_pred), ('pitch', self.pitch_pred), ('energy', self.energy_pred)):
            packs[name] = _accel.TransformerPack(mod.transformer, precise=(name == 'dur'))
            packs[name + '_emb'] = f(mod.embedding.weight)
            packs[name + '_lin'] = (f(mod.lin.weight), f(mod.lin.bias))
        return packs

    def _program(self, st: dict, key, builder):
        self._check_len(st['packs']['prenet'], key[2])       # the ACTUAL length; programs are built per capacity bucket
        return super()._program(st, key, builder)

    def _check_len(self, tp: '_accel.TransformerPack', n: int):
        if n > tp.max_len:
            raise RuntimeError(f'sequence length {n} exceeds PositionalEncoding.max_len={tp.max_len} '
                               '(same limit as the reference, common_layers.py:129)')

    def _token_program(self, st: dict, batch: int, T: int, predictors: bool,
                       mask_predictors: bool = False) -> eng.Program:
        packs, h = st['packs'], st['handle']
        geo = eng.Geometry(batch, T)
        prog = eng.Program(h, geo)
        tok = prog.dense('tok', (batch, -1), torch.int64)
        branches, heads = [], []

        def add_transformer(tag, tp, table, key_tok):
            E = tp.d_model
            opw, opdt, rmode = _accel.transformer_operand(tp)
            x0 = prog.buf(f'{tag}.x', opw * E, opdt)
            x0f = prog.buf(f'{tag}.x_full', E)
            out = prog.buf(f'{tag}.out', E)
            prog.call(lambda: h.posenc(tok, table, None, tp.pe, tp.scale, E, batch, geo.seq_len, geo.stride, x0,
                                       round_tf32=rmode, out_full=x0f),
                      label=f'{tag}.embed+posenc')                      # embedding + x + scale*pe (:143-145)
            branches.append(_accel.transformer_levels(prog, tag, tp, x0, out, key_tok, round_final=False, x_full=x0f))
            return out

        if predictors:
            for name in ('dur', 'pitch', 'energy'):
                # generate(): no mask (:174-180); forward(batch): the token pad mask (:132-135)
                out = add_transformer(name, packs[name], packs[name + '_emb'], tok if mask_predictors else None)
                heads.append((name, out))
        add_transformer('prenet', packs['prenet'], packs['embedding'], tok)            # key mask x == 0 (:200)
        eng.run_levels(prog, branches)
        for name, out in heads:
            w, b = packs[name + '_lin']
            y = prog.dense(f'{name}.y', (batch, -1))
            prog.call(lambda out=out, w=w, b=b, y=y, name=name: h.rowdot(
                out, out.shape[1], out.shape[1], w, b, 1, prog.alpha if name == 'dur' else 1.0,
                batch, geo.seq_len, geo.stride, y), label=f'{name}.lin')
        prog.alpha = 1.0
        return prog

    def _frame_program(self, st: dict, batch: int, M: int, masked: bool = False) -> eng.Program:
        """LR output -> postnet transformer -> lin -> [B, n_mels, M]  (fast_pitch.py:213-218).  generate(): no mask;
        `masked`: keys where the caller-filled prog.bufs['frame_tok'] is 0 are masked (forward(batch), :151-155)."""
        packs, h = st['packs'], st['handle']
        tp = packs['postnet']
        geo = eng.Geometry(batch, M)
        prog = eng.Program(h, geo)
        E = tp.d_model
        lr = prog.buf('lr', E)
        opw, opdt, rmode = _accel.transformer_operand(tp)
        x0 = prog.buf('postnet.x', opw * E, opdt)
        x0f = prog.buf('postnet.x_full', E)
        out = prog.buf('postnet.out', E)
        mel = prog.dense('mel', (batch, self.n_mels, -1))
        prog.call(lambda: h.posenc(None, None, lr, tp.pe, tp.scale, E, batch, geo.seq_len, geo.stride, x0, round_tf32=rmode, out_full=x0f),
                  label='postnet.posenc')
        frame_tok = prog.dense('frame_tok', (batch, -1), torch.int64) if masked else None
        eng.run_levels(prog, [_accel.transformer_levels(prog, 'postnet', tp, x0, out, frame_tok, x_full=x0f)])
        prog.label('lin')
        prog.gemm(out, packs['lin'], None, out_t=mel)
        prog.flush()
        prog.label('')
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
            prog = self._program(st, ('tok', B, T), lambda T: self._token_program(st, B, T, True))
            prog.bufs['tok'].copy_(x)
            prog.alpha = float(alpha)
            prog.run()
            dur_hat = prog.bufs['dur.y'].clone()
            pending = _accel.length_regulate_begin(st['handle'], dur_hat, apply_fallback=True)      # :176-177 (+ LR clamp)
            # the callbacks do not depend on the regulated length: they are enqueued while the scan result travels to the host
            pitch_hat = pitch_function(prog.bufs['pitch.y'].clone().unsqueeze(1))
            energy_hat = energy_function(prog.bufs['energy.y'].clone().unsqueeze(1))
            scan = pending.finish()
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, scan)

    def _forward_eval(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """forward(batch) in eval mode under no_grad on the sm_100a kernels (fast_pitch.py:122-165): predictors and
        prenet with the token pad mask, teacher-forced durations / pitch / energy, postnet with the frame mask
        m >= mel_len[b], outputs cut / padded to the target length."""
        x = batch['x']
        self._check_tokens(x)
        st = self._accel_state(x.device)
        B, T = x.shape
        tok = self._program(st, ('tok_masked', B, T), lambda T: self._token_program(st, B, T, True, mask_predictors=True))
        tok.bufs['tok'].copy_(x)
        tok.alpha = 1.0
        tok.run()
        dur_hat = tok.bufs['dur.y'].clone()
        pitch_hat = tok.bufs['pitch.y'].clone().unsqueeze(1)
        energy_hat = tok.bufs['energy.y'].clone().unsqueeze(1)
        dur = batch['dur'].to(device=x.device, dtype=torch.float32).contiguous().clone()
        cum, lens, M, _ = _accel.length_regulate(st['handle'], dur, apply_fallback=False)
        if M <= 0:
            raise RuntimeError('forward(): all durations are zero -- nothing to synthesise')
        prog = self._program(st, ('frame_masked', B, M), lambda M: self._frame_program(st, B, M, masked=True))
        mel_len = batch['mel_len'].to(x.device)
        prog.bufs['frame_tok'].copy_((torch.arange(M, device=x.device)[None, :] < mel_len[:, None]).to(torch.int64))
        out = _accel.gather_and_run_frames(st['handle'], st['packs'], tok, prog, tok.bufs['prenet.out'], None, 0, dur,
                                           batch['pitch'].to(x.device), batch['energy'].to(x.device), cum, lens, M,
                                           float(self.pitch_strength), float(self.energy_strength), round_tf32=False)
        mel = self.pad(out['mel'], batch['mel'].size(2))
        return {'mel': mel, 'mel_post': mel.clone(), 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}

    def _generate_mel(self, x: torch.Tensor, dur_hat: torch.Tensor, pitch_hat: torch.Tensor,
                      energy_hat: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Teacher-forced entry (fast_pitch.py:194-221)."""
        self._check_tokens(x)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('prenet', B, T), lambda T: self._token_program(st, B, T, False))
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
        pre = tok_prog.bufs['prenet.out']
        prog = self._program(st, ('frame', B, M), lambda M: self._frame_program(st, B, M))
        return _accel.gather_and_run_frames(h, packs, tok_prog, prog, pre, None, 0, dur_hat, pitch_hat, energy_hat,
                                            cum, lens, M, float(self.pitch_strength), float(self.energy_strength),
                                            round_tf32=False)

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> 'FastPitch':
        kwargs = config['fast_pitch']['model']
        kwargs['num_chars'] = NUM_CHARS
        kwargs['n_mels'] = config['dsp']['num_mels']
        return cls(**kwargs)

    @classmethod
    def from_checkpoint(cls, path: Union[Path, str]) -> 'FastPitch':
        ckpt = torch.load(path, map_location=torch.device('cpu'))
        model = cls.from_config(ckpt['config'])
        model.load_state_dict(ckpt['model'])
        return model
