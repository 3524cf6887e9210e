"""MultiFastPitch (multi-speaker FFT-block variant) with the single-pass inference path on sm_100a kernels.

Drop-in for the reference's `models/multi_fast_pitch.py`: same constructor kwargs (:94-131), same state_dict, same
`generate(x, speaker_emb, alpha, pitch_function, energy_function)` surface (:233-237).  SURVEY.md §8f "next" row.

What differs from FastPitch: every predictor's transformer input is the concatenation
[embedding(x), conditional_embedding(pitch_cond), speaker_emb] (:43-48, :76-87; widths 384 / 392), the main branch
concatenates the speaker vector BEFORE the prenet (:277-283; width 512), and generation is two-phase: the
pitch-condition classifier runs first, its arg-max feeds dur_pred and pitch_pred (:242-249).

Reference semantics kept on purpose: predictors run WITHOUT a key-padding mask in generate(); the prenet masks keys
where the token id is 0 (:275); the postnet attends over all M frames (:295); 'mel' and 'mel_post' are the same tensor.
Documented deviation (as in MultiForwardTacotron): the reference's arg-max line (:243) is only well-formed for batch
1; here it is taken per (b,t) over the class axis, which is the same thing at batch 1; 'pitch_cond' is [B,T].
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


def _with_speaker(e: torch.Tensor, speaker_emb: torch.Tensor) -> torch.Tensor:
    return torch.cat([e, speaker_emb[:, None, :].repeat(1, e.shape[1], 1)], dim=2)


class SeriesPredictor(nn.Module):
    """[embedding(x), speaker] -> ForwardTransformer -> Linear(out_dim) -> /alpha  (multi_fast_pitch.py:14-50)."""

    def __init__(self, num_chars: int, d_model: int, n_heads: int, d_fft: int, layers: int, conv1_kernel: int,
                 conv2_kernel: int, speaker_emb_dims: int, dropout=0.1, out_dim: int = 1):
        super().__init__()
        self.embedding = nn.Embedding(num_chars, d_model)
        self.transformer = ForwardTransformer(heads=n_heads, dropout=dropout, d_model=d_model + speaker_emb_dims,
                                              d_fft=d_fft, conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel,
                                              layers=layers)
        self.lin = nn.Linear(d_model + speaker_emb_dims, out_dim)

    def forward(self, x: torch.Tensor, speaker_emb: torch.Tensor, src_pad_mask: Optional[torch.Tensor] = None,
                alpha: float = 1.0) -> torch.Tensor:
        y = self.transformer(_with_speaker(self.embedding(x), speaker_emb), src_pad_mask=src_pad_mask)
        return self.lin(y) / alpha


class ConditionalSeriesPredictor(nn.Module):
    """[embedding(x), conditional_embedding(x_cond), speaker] -> ForwardTransformer -> Linear(1) -> /alpha (:53-91)."""

    def __init__(self, num_chars: int, d_model: int, n_heads: int, d_fft: int, layers: int, conv1_kernel: int,
                 conv2_kernel: int, speaker_emb_dims: int, cond_emb_size: int = 4, cond_emb_dims: int = 8,
                 dropout=0.1):
        super().__init__()
        self.embedding = nn.Embedding(num_chars, d_model)
        self.conditional_embedding = nn.Embedding(cond_emb_size, cond_emb_dims)
        width = d_model + speaker_emb_dims + cond_emb_dims
        self.transformer = ForwardTransformer(heads=n_heads, dropout=dropout, d_model=width, d_fft=d_fft,
                                              conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel, layers=layers)
        self.lin = nn.Linear(width, 1)

    def forward(self, x: torch.Tensor, x_cond: torch.Tensor, speaker_emb: torch.Tensor,
                src_pad_mask: Optional[torch.Tensor] = None, alpha: float = 1.0) -> torch.Tensor:
        e = torch.cat([self.embedding(x), self.conditional_embedding(x_cond)], dim=2)
        y = self.transformer(_with_speaker(e, speaker_emb), src_pad_mask=src_pad_mask)
        return self.lin(y) / alpha


class MultiFastPitch(nn.Module, _accel.AcceleratedMixin):

    def __init__(self,
                 num_chars: int,
                 durpred_dropout: float, durpred_d_model: int, durpred_n_heads: int, durpred_layers: int,
                 durpred_d_fft: int,
                 pitch_dropout: float, pitch_d_model: int, pitch_n_heads: int, pitch_layers: int, pitch_d_fft: int,
                 energy_dropout: float, energy_d_model: int, energy_n_heads: int, energy_layers: int,
                 energy_d_fft: int,
                 pitch_cond_d_model: int, pitch_cond_n_heads: int, pitch_cond_layers: int, pitch_cond_d_fft: int,
                 pitch_cond_dropout: float, pitch_cond_output_dims: int,
                 pitch_strength: float, energy_strength: float,
                 d_model: int, conv1_kernel: int, conv2_kernel: int,
                 prenet_layers: int, prenet_heads: int, prenet_fft: int, prenet_dropout: float,
                 postnet_layers: int, postnet_heads: int, postnet_fft: int, postnet_dropout: float,
                 n_mels: int, speaker_emb_dims: int, padding_value=-11.5129):
        super().__init__()
        self.padding_value = padding_value
        self.n_mels = n_mels
        self.speaker_emb_dims = speaker_emb_dims
        self.lr = LengthRegulator()
        common = dict(num_chars=num_chars, conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel,
                      speaker_emb_dims=speaker_emb_dims)
        self.dur_pred = ConditionalSeriesPredictor(d_model=durpred_d_model, n_heads=durpred_n_heads,
                                                   layers=durpred_layers, d_fft=durpred_d_fft,
                                                   dropout=durpred_dropout, **common)
        self.pitch_pred = ConditionalSeriesPredictor(d_model=pitch_d_model, n_heads=pitch_n_heads,
                                                     layers=pitch_layers, d_fft=pitch_d_fft, dropout=pitch_dropout,
                                                     **common)
        self.pitch_cond_pred = SeriesPredictor(d_model=pitch_cond_d_model, n_heads=pitch_cond_n_heads,
                                               layers=pitch_cond_layers, d_fft=pitch_cond_d_fft,
                                               dropout=pitch_cond_dropout, out_dim=pitch_cond_output_dims, **common)
        self.energy_pred = SeriesPredictor(d_model=energy_d_model, n_heads=energy_n_heads, layers=energy_layers,
                                           d_fft=energy_d_fft, dropout=energy_dropout, **common)
        self.embedding = nn.Embedding(num_embeddings=num_chars, embedding_dim=d_model)
        width = d_model + speaker_emb_dims
        kern = dict(conv1_kernel=conv1_kernel, conv2_kernel=conv2_kernel)
        self.prenet = ForwardTransformer(heads=prenet_heads, dropout=prenet_dropout, d_model=width,
                                         d_fft=prenet_fft, layers=prenet_layers, **kern)
        self.postnet = ForwardTransformer(heads=postnet_heads, dropout=postnet_dropout, d_model=width,
                                          d_fft=postnet_fft, layers=postnet_layers, **kern)
        self.lin = nn.Linear(width, n_mels)
        self.register_buffer('step', torch.zeros(1, dtype=torch.long))
        self.pitch_strength = pitch_strength
        self.energy_strength = energy_strength
        self.pitch_proj = nn.Conv1d(1, width, kernel_size=3, padding=1)
        self.energy_proj = nn.Conv1d(1, width, kernel_size=3, padding=1)

    def __repr__(self):
        return f'MultiFastPitch, num params: {sum(int(np.prod(p.size())) for p in self.parameters())}'

    def get_step(self) -> int:
        return self.step.data.item()

    # ------------------------------------------------------------------ training path (stock torch)
    def forward(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        x, semb, mel_lens, pitch_cond = batch['x'], batch['speaker_emb'], batch['mel_len'], batch['pitch_cond']
        target_len = batch['mel'].size(2)
        if self.training:
            self.step += 1
        if not self.training and x.is_cuda and not torch.is_grad_enabled():
            return self._forward_eval(batch)               # the GTA-dump / evaluation use (train_forward.py:33-51)
        tok_mask = make_token_len_mask(x.transpose(0, 1))
        dur_hat = self.dur_pred(x, x_cond=pitch_cond, speaker_emb=semb, src_pad_mask=tok_mask).squeeze(-1)
        pitch_hat = self.pitch_pred(x, x_cond=pitch_cond, speaker_emb=semb, src_pad_mask=tok_mask).transpose(1, 2)
        pitch_cond_hat = self.pitch_cond_pred(x, speaker_emb=semb, src_pad_mask=tok_mask)
        energy_hat = self.energy_pred(x, speaker_emb=semb, src_pad_mask=tok_mask).transpose(1, 2)
        h = self.prenet(_with_speaker(self.embedding(x), semb), src_pad_mask=tok_mask)
        h = h + self.pitch_proj(batch['pitch'].unsqueeze(1)).transpose(1, 2) * self.pitch_strength
        h = h + self.energy_proj(batch['energy'].unsqueeze(1)).transpose(1, 2) * self.energy_strength
        h = self.lr(h, batch['dur'])
        frame_mask = torch.arange(h.size(1), device=h.device)[None, :] >= mel_lens.to(h.device)[:, None]
        h = self.postnet(h, src_pad_mask=frame_mask)
        mel = self.pad(self.lin(h).transpose(1, 2), target_len)
        return {'mel': mel, 'mel_post': mel.clone(), 'pitch_cond': pitch_cond_hat, 'dur': dur_hat,
                'pitch': pitch_hat, 'energy': energy_hat}

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
        for name, mod in (('dur', self.dur_pred), ('pitch', self.pitch_pred), ('energy', self.energy_pred),
                          ('pitch_cond', self.pitch_cond_pred)):
            # durations and the pitch-condition arg-max are rounded to integers: split-precision predictors
            packs[name] = _accel.TransformerPack(mod.transformer, precise=name in ('dur', 'pitch_cond'))
            packs[name + '_emb'] = f(mod.embedding.weight)
            if hasattr(mod, 'conditional_embedding'):
                packs[name + '_cemb'] = f(mod.conditional_embedding.weight)
            packs[name + '_lin'] = (f(mod.lin.weight), f(mod.lin.bias))
        return packs

    def _program(self, st: dict, key, builder):
        self._check_len(st['packs']['prenet'], key[2])       # the ACTUAL length; programs are built per capacity bucket
        return super()._program(st, key, builder)

    def _check_len(self, tp: '_accel.TransformerPack', n: int):
        if n > tp.max_len:
            raise RuntimeError(f'sequence length {n} exceeds PositionalEncoding.max_len={tp.max_len} '
                               '(same limit as the reference, common_layers.py:129)')

    def _token_program(self, st: dict, batch: int, T: int, predictors: bool, mask_predictors: bool = False,
                       teacher_cond: bool = False) -> eng.Program:
        """Phase 1: pitch_cond_pred, energy_pred and the prenet transformer (level interleaved), arg-max of the pitch
        condition (:242-243).  Phase 2: dur_pred and pitch_pred, which consume it (:244,248).  Eval-mode forward(batch)
        (:205-210): `mask_predictors` gives every predictor the token pad mask, `teacher_cond` takes the condition from
        the caller-filled 'pitch_cond' buffer instead of the arg-max."""
        packs, h = st['packs'], st['handle']
        geo = eng.Geometry(batch, T)
        prog = eng.Program(h, geo)
        tok = prog.dense('tok', (batch, -1), torch.int64)
        semb = prog.dense('semb', (batch, self.speaker_emb_dims))
        pcond = prog.dense('pitch_cond', (batch, -1), torch.int64)

        def add_transformer(tag, tp, table, ctable, key_tok):
            """concat [table[tok], ctable[pcond], semb] -> + scale*pe -> transformer levels."""
            E = tp.d_model
            cat = prog.buf(f'{tag}.cat', E)
            opw, opdt, rmode = _accel.transformer_operand(tp)
            x0 = prog.buf(f'{tag}.x', opw * E, opdt)
            x0f = prog.buf(f'{tag}.x_full', E)
            out = prog.buf(f'{tag}.out', E)
            prog.call(lambda: h.embed(tok, table, cat, E, batch, geo.seq_len, geo.stride,
                                      tok2=pcond if ctable is not None else None, table2=ctable, extra=semb,
                                      round_tf32=False), label=f'{tag}.embed')
            prog.call(lambda: h.posenc(None, None, cat, tp.pe, tp.scale, E, batch, geo.seq_len, geo.stride, x0,
                                       round_tf32=rmode, out_full=x0f),
                      label=f'{tag}.posenc')
            return out, _accel.transformer_levels(prog, tag, tp, x0, out, key_tok, round_final=False, x_full=x0f)

        def add_head(name, out):
            w, b = packs[name + '_lin']
            n_out = w.shape[0]
            y = prog.dense(f'{name}.y', (batch, -1, n_out))
            alpha_of = (lambda: prog.alpha) if name in ('dur', 'pitch_cond') else (lambda: 1.0)
            prog.call(lambda: h.rowdot(out, out.shape[1], out.shape[1], w, b, n_out, alpha_of(), batch, geo.seq_len,
                                       geo.stride, y), label=f'{name}.lin')
            return y

        branches, outs = [], {}
        if predictors:
            for name in ('pitch_cond', 'energy'):
                outs[name], lv = add_transformer(name, packs[name], packs[name + '_emb'], None,
                                                 tok if mask_predictors else None)
                branches.append(lv)
        _, lv = add_transformer('prenet', packs['prenet'], packs['embedding'], None, tok)   # key mask x == 0 (:275)
        branches.append(lv)
        eng.run_levels(prog, branches)
        if predictors:
            logits = add_head('pitch_cond', outs['pitch_cond'])
            add_head('energy', outs['energy'])
            if not teacher_cond:
                prog.call(lambda: h.argmax(logits, batch * geo.seq_len, logits.shape[2], pcond), label='pitch_cond.argmax')
            branches2 = []
            for name in ('dur', 'pitch'):
                outs[name], lv = add_transformer(name, packs[name], packs[name + '_emb'], packs[name + '_cemb'],
                                                 tok if mask_predictors else None)
                branches2.append(lv)
            eng.run_levels(prog, branches2)
            add_head('dur', outs['dur'])
            add_head('pitch', outs['pitch'])
        prog.alpha = 1.0
        return prog

    def _frame_program(self, st: dict, batch: int, M: int, masked: bool = False) -> eng.Program:
        """LR output -> postnet transformer -> lin -> [B, n_mels, M]  (:293-298).  generate(): no mask; `masked`: keys
        where the caller-filled prog.bufs['frame_tok'] is 0 are masked (forward(batch), :230-234)."""
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
            pitch_cond_hat = prog.bufs['pitch_cond'].clone()                    # [B,T] int64
            dur_hat = prog.bufs['dur.y'].reshape(B, T).clone()
            pending = _accel.length_regulate_begin(st['handle'], dur_hat, apply_fallback=True)      # :246-247 (+ LR clamp)
            # the callbacks do not depend on the regulated length: they are enqueued while the scan result travels to the host
            pitch_hat = pitch_function(prog.bufs['pitch.y'].reshape(B, 1, T).clone())
            energy_hat = energy_function(prog.bufs['energy.y'].reshape(B, 1, T).clone())
            scan = pending.finish()
            return self._mel_from_prenet(st, prog, dur_hat, pitch_hat, energy_hat, pitch_cond_hat, scan)

    def _forward_eval(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """forward(batch) in eval mode under no_grad on the sm_100a kernels (multi_fast_pitch.py:192-245): predictors
        and prenet with the token pad mask, dur / pitch conditioned on the target pitch condition, teacher-forced
        durations / pitch / energy, postnet with the frame mask m >= mel_len[b], outputs cut / padded to the target."""
        x, semb = batch['x'], batch['speaker_emb']
        self._check_tokens(x)
        self._check_speaker(x, semb, self.speaker_emb_dims)
        st = self._accel_state(x.device)
        B, T = x.shape
        tok = self._program(st, ('tok_masked', B, T),
                            lambda T: self._token_program(st, B, T, True, mask_predictors=True, teacher_cond=True))
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
        return {'mel': mel, 'mel_post': mel.clone(), 'pitch_cond': pitch_cond_hat, 'dur': dur_hat,
                'pitch': pitch_hat, 'energy': energy_hat}

    def _generate_mel(self, x: torch.Tensor, speaker_emb: torch.Tensor, dur_hat: torch.Tensor,
                      pitch_hat: torch.Tensor, pitch_cond_hat: torch.Tensor,
                      energy_hat: torch.Tensor) -> Dict[str, torch.Tensor]:
        """Teacher-forced entry (multi_fast_pitch.py:266-305)."""
        self._check_tokens(x)
        self._check_speaker(x, speaker_emb, self.speaker_emb_dims)
        with torch.no_grad():
            st = self._accel_state(x.device)
            B, T = x.shape
            prog = self._program(st, ('prenet', B, T), lambda T: self._token_program(st, B, T, False))
            prog.bufs['tok'].copy_(x)
            prog.bufs['semb'].copy_(speaker_emb.to(torch.float32))
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
        pre = tok_prog.bufs['prenet.out']
        prog = self._program(st, ('frame', B, M), lambda M: self._frame_program(st, B, M))
        out = _accel.gather_and_run_frames(h, packs, tok_prog, prog, pre, None, 0, dur_hat, pitch_hat, energy_hat,
                                           cum, lens, M, float(self.pitch_strength), float(self.energy_strength),
                                           round_tf32=False)
        out['pitch_cond'] = pitch_cond_hat.reshape(B, T)
        return out

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> 'MultiFastPitch':
        kwargs = config['multi_fast_pitch']['model']
        kwargs['num_chars'] = NUM_CHARS
        kwargs['n_mels'] = config['dsp']['num_mels']
        return cls(**kwargs)

    @classmethod
    def from_checkpoint(cls, path: Union[Path, str]) -> 'MultiFastPitch':
        ckpt = torch.load(path, map_location=torch.device('cpu'))
        model = cls.from_config(ckpt['config'])
        model.load_state_dict(ckpt['model'])
        return model
