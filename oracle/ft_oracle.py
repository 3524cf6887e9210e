"""CPU oracle for the mel-synthesis forward path  --  TEST INFRASTRUCTURE ONLY.

This file is the checker, not the product: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / `--impl reference` legs may import it.  Nothing in
forwardtacotron_b200/ imports from oracle/; the product path raises when its
CUDA library is missing instead of falling back to this.

What it restates.  The reference (ziyaad30/ForwardTacotron) is pure
Python/PyTorch; its arithmetic lives in PyTorch itself (ATen/oneDNN on CPU), an
un-vendored, un-pinned third-party dependency (requirements.txt does not list
torch; this image has torch 2.11.0).  Each function below restates one reference
function as a pure function of (state_dict, inputs) built from the published
layer definitions (Conv1d cross-correlation with zero padding, eval-mode
BatchNorm1d, GRU/LSTM gate equations from the torch.nn docs, LayerNorm,
scaled-dot-product attention) and cites the reference file:line it follows.
The integer/index part (LengthRegulator) is restated in numpy and is bit-exact.

Parity pinning.  The reference's own tests hold no numeric golden for this path
(tests/test_forward_tacotron.py:40-46 pins shapes/keys only), so the oracle is
pinned against outputs of the reference itself: oracle/make_golden.py imports
/root/reference, loads key-derived synthetic weights
(forwardtacotron_b200/synth.py), runs the reference's generate() and commits
the outputs under tests/golden/; tests/test_oracle_golden.py checks every
function here against those files.  Status: PINNED.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Mapping, Optional

import numpy as np
import torch
import torch.nn.functional as F

SD = Mapping[str, torch.Tensor]
Identity = lambda t: t  # noqa: E731

# Optional operand-rounding hook (precision studies: emulate tf32/bf16 GEMM
# operands on the CPU to size the error budget per kernel).
_QUANT: Callable[[torch.Tensor], torch.Tensor] = Identity


_QUANT_REC: Optional[Callable[[torch.Tensor], torch.Tensor]] = None   # None: follow _QUANT


def set_operand_rounding(fn: Optional[Callable[[torch.Tensor], torch.Tensor]],
                         recurrent: Optional[Callable[[torch.Tensor], torch.Tensor]] = None) -> None:
    """`fn` rounds the operands of every feed-forward contraction; `recurrent` (default: same
    as fn) those of the h @ W_hh^T products inside the GRU/LSTM time loops."""
    global _QUANT, _QUANT_REC
    _QUANT = fn if fn is not None else Identity
    _QUANT_REC = recurrent


def _qrec(t: torch.Tensor) -> torch.Tensor:
    return (_QUANT_REC or _QUANT)(t)


def round_tf32(t: torch.Tensor) -> torch.Tensor:
    """Round-to-nearest-even onto the 10-bit tf32 mantissa (what cvt.rna.tf32 does,
    up to tie direction)."""
    i = t.contiguous().view(torch.int32)
    bias = ((i >> 13) & 1) + 0x0FFF
    return ((i + bias) & ~0x1FFF).view(torch.float32)


def round_bf16(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


# --------------------------------------------------------------------------
# LengthRegulator  (models/common_layers.py:12-24)  -- integer work, bit exact
# --------------------------------------------------------------------------

def length_regulate_np(x: np.ndarray, dur: np.ndarray):
    """x [B,T,C] fp32, dur [B,T] fp32  ->  (out [B,M,C], idx [B,M] int32, lens [B] int32, dur_clamped).

    common_layers.py:18   dur[dur < 0] = 0              (in place in the reference)
    common_layers.py:21   rep = (dur[i] + 0.5).long()    fp32 add, then trunc toward zero
    common_layers.py:21   repeat_interleave(x[i], rep)   row m copies x[i, idx[m]]
    common_layers.py:23   pad_sequence(..., 0.)          rows >= len_b are 0.0, M = max_b len_b
    idx[b, m] = -1 marks a padded row.
    """
    x = np.ascontiguousarray(x, dtype=np.float32)
    d = np.array(dur, dtype=np.float32, copy=True)
    d[d < 0] = np.float32(0.0)
    rep = (d + np.float32(0.5)).astype(np.float32)     # fp32 add (round-to-nearest-even)
    rep = np.trunc(rep).astype(np.int64)
    lens = rep.sum(axis=1).astype(np.int32)
    B, T, C = x.shape
    M = int(lens.max()) if B else 0
    out = np.zeros((B, M, C), dtype=np.float32)
    idx = np.full((B, M), -1, dtype=np.int32)
    for b in range(B):
        src = np.repeat(np.arange(T, dtype=np.int32), rep[b])
        idx[b, :src.shape[0]] = src
        out[b, :src.shape[0]] = x[b, src]
    return out, idx, lens, d


def length_regulate(x: torch.Tensor, dur: torch.Tensor) -> torch.Tensor:
    out, _, _, _ = length_regulate_np(x.numpy(), dur.numpy())
    return torch.from_numpy(out)


# --------------------------------------------------------------------------
# Building blocks
# --------------------------------------------------------------------------

def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    return F.linear(_QUANT(x), _QUANT(w), b)


def batch_norm_conv(sd: SD, p: str, x: torch.Tensor, relu: bool) -> torch.Tensor:
    """BatchNormConv.forward, common_layers.py:54-57: conv(pad=k//2, no bias) -> ReLU -> BN(eval).
    x [B,Cin,L] -> [B,Cout,L] (L+1 for even k; callers trim, common_layers.py:99)."""
    w = sd[p + '.conv.weight']
    k = w.shape[2]
    y = F.conv1d(_QUANT(x), _QUANT(w), None, stride=1, padding=k // 2)
    if relu:
        y = F.relu(y)
    mean, var = sd[p + '.bnorm.running_mean'], sd[p + '.bnorm.running_var']
    gamma, beta = sd[p + '.bnorm.weight'], sd[p + '.bnorm.bias']
    # nn.BatchNorm1d in eval mode (common_layers.py:57): the same fused ATen op the reference's module calls
    return F.batch_norm(y, mean, var, gamma, beta, False, 0.0, 1e-5)


def highway(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """HighwayNetwork.forward, common_layers.py:35-40."""
    x1 = linear(x, sd[p + '.W1.weight'], sd[p + '.W1.bias'])
    x2 = linear(x, sd[p + '.W2.weight'], sd[p + '.W2.bias'])
    g = torch.sigmoid(x2)
    return g * F.relu(x1) + (1.0 - g) * x


def gru_direction(x: torch.Tensor, w_ih, w_hh, b_ih, b_hh, reverse: bool) -> torch.Tensor:
    """One direction of nn.GRU, gate order (r,z,n), h0 = 0, over the full padded length.
       r = s(W_ir x + b_ir + W_hr h + b_hr);  z likewise;
       n = tanh(W_in x + b_in + r*(W_hn h + b_hn));  h' = (1-z)*n + z*h      (torch.nn.GRU docs)."""
    B, L, _ = x.shape
    H = w_hh.shape[1]
    gi = linear(x, w_ih, b_ih)                       # [B,L,3H]
    h = x.new_zeros(B, H)
    out = x.new_zeros(B, L, H)
    wq = _qrec(w_hh)
    steps = range(L - 1, -1, -1) if reverse else range(L)
    for t in steps:
        gh = F.linear(_qrec(h), wq, b_hh)
        i_r, i_z, i_n = gi[:, t].chunk(3, dim=1)
        h_r, h_z, h_n = gh.chunk(3, dim=1)
        r = torch.sigmoid(i_r + h_r)
        z = torch.sigmoid(i_z + h_z)
        n = torch.tanh(i_n + r * h_n)
        h = (1.0 - z) * n + z * h
        out[:, t] = h
    return out


def lstm_direction(x: torch.Tensor, w_ih, w_hh, b_ih, b_hh, reverse: bool) -> torch.Tensor:
    """One direction of nn.LSTM, gate order (i,f,g,o), h0 = c0 = 0 (torch.nn.LSTM docs)."""
    B, L, _ = x.shape
    H = w_hh.shape[1]
    gi = linear(x, w_ih, b_ih)
    h = x.new_zeros(B, H)
    c = x.new_zeros(B, H)
    out = x.new_zeros(B, L, H)
    wq = _qrec(w_hh)
    steps = range(L - 1, -1, -1) if reverse else range(L)
    for t in steps:
        g = gi[:, t] + F.linear(_qrec(h), wq, b_hh)
        gi_, gf_, gg_, go_ = g.chunk(4, dim=1)
        c = torch.sigmoid(gf_) * c + torch.sigmoid(gi_) * torch.tanh(gg_)
        h = torch.sigmoid(go_) * torch.tanh(c)
        out[:, t] = h
    return out


_FAST_RNN = False


def use_library_rnn(flag: bool) -> None:
    """Route bi_rnn through torch's fused CPU GRU/LSTM (the very kernel the reference
    calls) instead of the explicit per-step loops -- used for the timed CPU baseline."""
    global _FAST_RNN
    _FAST_RNN = bool(flag)


def bi_rnn(sd: SD, p: str, x: torch.Tensor, kind: str) -> torch.Tensor:
    """Bidirectional batch_first GRU/LSTM over the full padded length (no packing in
    generate(): common_layers.py:123, forward_tacotron.py:37,224).  out = [fwd | bwd]."""
    names = ['weight_ih_l0', 'weight_hh_l0', 'bias_ih_l0', 'bias_hh_l0']
    if _FAST_RNN and _QUANT is Identity and _QUANT_REC is None:
        flat = [sd[f'{p}.{n}{sfx}'] for sfx in ('', '_reverse') for n in names]
        B = x.shape[0]
        H = flat[1].shape[1]
        h0 = x.new_zeros(2, B, H)
        if kind == 'gru':
            return torch._VF.gru(x, h0, flat, True, 1, 0.0, False, True, True)[0]
        return torch._VF.lstm(x, (h0, h0.clone()), flat, True, 1, 0.0, False, True, True)[0]
    step = gru_direction if kind == 'gru' else lstm_direction
    outs = []
    for sfx, rev in (('', False), ('_reverse', True)):
        outs.append(step(x, *[sd[f'{p}.{n}{sfx}'] for n in names], reverse=rev))
    return torch.cat(outs, dim=2)


def cbhg(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """CBHG.forward (eval), common_layers.py:91-124.  x [B,Cin,L] -> [B,L,2*channels]."""
    L = x.shape[-1]
    bank = []
    k = 0
    while f'{p}.conv1d_bank.{k}.conv.weight' in sd:
        bank.append(batch_norm_conv(sd, f'{p}.conv1d_bank.{k}', x, True)[:, :, :L])   # :97-99
        k += 1
    y = torch.cat(bank, dim=1)                                                         # :102
    y = F.max_pool1d(y, kernel_size=2, stride=1, padding=1)[:, :, :L]                  # :105
    y = batch_norm_conv(sd, p + '.conv_project1', y, True)                             # :109
    y = batch_norm_conv(sd, p + '.conv_project2', y, False)                            # :111
    y = y + x                                                                          # :114
    y = y.transpose(1, 2)
    y = linear(y, sd[p + '.pre_highway.weight'])                                       # :118
    h = 0
    while f'{p}.highways.{h}.W1.weight' in sd:
        y = highway(sd, f'{p}.highways.{h}', y)                                        # :119-120
        h += 1
    return bi_rnn(sd, p + '.rnn', y, 'gru')                                            # :123


def series_predictor(sd: SD, p: str, x_emb: torch.Tensor, alpha: float = 1.0) -> torch.Tensor:
    """conv stack -> BiGRU -> Linear -> /alpha of SeriesPredictor.forward
    (forward_tacotron.py:32-39; multi_forward_tacotron.py:42-50,83-93).
    x_emb [B,T,E] is the (concatenated) embedding input."""
    y = x_emb.transpose(1, 2)
    for i in range(3):
        y = batch_norm_conv(sd, f'{p}.convs.{i}', y, True)
    y = y.transpose(1, 2)
    y = bi_rnn(sd, p + '.rnn', y, 'gru')
    y = linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias'])
    return y / alpha


def cond_projection(sd: SD, p: str, series: torch.Tensor) -> torch.Tensor:
    """pitch_proj / energy_proj: Conv1d(1->C, k=3, pad=1, bias) then transpose
    (forward_tacotron.py:214-215).  series [B,1,T] -> [B,T,C].  Not a GEMM on the GPU
    (3-tap FIR), so no operand rounding here."""
    y = F.conv1d(series, sd[p + '.weight'], sd[p + '.bias'], padding=1)
    return y.transpose(1, 2)


def apply_fallback(dur_hat: torch.Tensor) -> torch.Tensor:
    """forward_tacotron.py:176-177: batch-global `sum(dur.long()) <= 0` -> fill 2.0."""
    if int(torch.sum(dur_hat.long())) <= 0:
        dur_hat = torch.full_like(dur_hat, 2.0)
    return dur_hat


# --------------------------------------------------------------------------
# ForwardTacotron   (models/forward_tacotron.py)
# --------------------------------------------------------------------------

@torch.no_grad()
def ft_generate_mel(sd: SD, x: torch.Tensor, dur_hat, pitch_hat, energy_hat,
                    pitch_strength: float = 1.0, energy_strength: float = 1.0,
                    taps: Optional[dict] = None) -> Dict[str, torch.Tensor]:
    """ForwardTacotron._generate_mel, forward_tacotron.py:205-234."""
    e = F.embedding(x, sd['embedding.weight']).transpose(1, 2)                 # :210-211
    h = cbhg(sd, 'prenet', e)                                                  # :212
    if taps is not None:
        taps['prenet'] = h
    h = h + cond_projection(sd, 'pitch_proj', pitch_hat) * pitch_strength      # :214-216
    h = h + cond_projection(sd, 'energy_proj', energy_hat) * energy_strength   # :218-220
    if h.is_cuda:
        # bench.py --impl eager-gpu only: the reference's own torch formulation of the LengthRegulator
        # (common_layers.py:17-24: clamp, (dur + 0.5).long(), repeat_interleave per sentence, pad_sequence)
        d = dur_hat.clone()
        d[d < 0] = 0
        rep = (d + 0.5).long()
        h = torch.nn.utils.rnn.pad_sequence([torch.repeat_interleave(h[i], rep[i], dim=0) for i in range(h.shape[0])],
                                            batch_first=True)
        lens_t, dur_t = rep.sum(dim=1), d
    else:
        expanded, idx, lens, dur_clamped = length_regulate_np(h.numpy(), dur_hat.numpy())   # :222
        h = torch.from_numpy(expanded)
        lens_t, dur_t = torch.from_numpy(lens), torch.from_numpy(dur_clamped)
        if taps is not None:
            taps['idx'] = torch.from_numpy(idx)
    if taps is not None:
        taps['lr'] = h
    h = bi_rnn(sd, 'lstm', h, 'lstm')                                          # :224
    if taps is not None:
        taps['lstm'] = h
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)          # :226-227
    post = cbhg(sd, 'postnet', mel)                                            # :229
    mel_post = linear(post, sd['post_proj.weight']).transpose(1, 2)            # :230-231
    return {'mel': mel, 'mel_post': mel_post, 'dur': dur_t,
            'pitch': pitch_hat, 'energy': energy_hat, 'mel_lens': lens_t}


@torch.no_grad()
def ft_generate(sd: SD, x: torch.Tensor, alpha: float = 1.0,
                pitch_function=Identity, energy_function=Identity,
                pitch_strength: float = 1.0, energy_strength: float = 1.0,
                taps: Optional[dict] = None) -> Dict[str, torch.Tensor]:
    """ForwardTacotron.generate, forward_tacotron.py:167-184."""
    def pred(p, a=1.0):
        return series_predictor(sd, p, F.embedding(x, sd[p + '.embedding.weight']), a)
    dur_hat = apply_fallback(pred('dur_pred', alpha).squeeze(2))
    pitch_hat = pitch_function(pred('pitch_pred').transpose(1, 2))
    energy_hat = energy_function(pred('energy_pred').transpose(1, 2))
    if taps is not None:
        taps['dur_raw'] = dur_hat.clone()
    return ft_generate_mel(sd, x, dur_hat, pitch_hat, energy_hat,
                           pitch_strength, energy_strength, taps)


@torch.no_grad()
def packed_bi_rnn(sd: SD, p: str, x: torch.Tensor, lens: torch.Tensor, kind: str, pad_value: float) -> torch.Tensor:
    """pack_padded_sequence -> bidirectional RNN -> pad_packed_sequence(padding_value) (forward_tacotron.py:147-152):
    every sentence runs over its own `lens[b]` steps (the backward direction starts at lens[b]-1 from a zero state);
    the result is padded to max(lens) with `pad_value`."""
    B, H2 = x.shape[0], 2 * sd[p + '.weight_hh_l0'].shape[1]
    M = int(lens.max())
    out = torch.full((B, M, H2), float(pad_value), dtype=torch.float32)
    for b in range(B):
        n = int(lens[b])
        out[b, :n] = bi_rnn(sd, p, x[b:b + 1, :n], kind)[0]
    return out


@torch.no_grad()
def ft_forward(sd: SD, batch: Dict[str, torch.Tensor], pitch_strength: float = 1.0, energy_strength: float = 1.0,
               padding_value: float = -11.5129) -> Dict[str, torch.Tensor]:
    """ForwardTacotron.forward(batch) in eval mode (forward_tacotron.py:118-165): teacher-forced durations / pitch /
    energy, packed-sequence BiLSTM over `mel_len`, outputs padded to the target mel length with `padding_value`."""
    x = batch['x']
    def pred(p):
        return series_predictor(sd, p, F.embedding(x, sd[p + '.embedding.weight']))
    dur_hat = pred('dur_pred').squeeze(-1)
    pitch_hat = pred('pitch_pred').transpose(1, 2)
    energy_hat = pred('energy_pred').transpose(1, 2)
    h = cbhg(sd, 'prenet', F.embedding(x, sd['embedding.weight']).transpose(1, 2))
    h = h + cond_projection(sd, 'pitch_proj', batch['pitch'].unsqueeze(1)) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', batch['energy'].unsqueeze(1)) * energy_strength
    expanded, _, _, _ = length_regulate_np(h.numpy(), batch['dur'].to(torch.float32).numpy().copy())
    h = packed_bi_rnn(sd, 'lstm', torch.from_numpy(expanded), batch['mel_len'], 'lstm', padding_value)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    mel_post = linear(cbhg(sd, 'postnet', mel), sd['post_proj.weight']).transpose(1, 2)
    target = batch['mel'].shape[2]
    def pad(t):
        t = t[:, :, :target]
        return F.pad(t, [0, target - t.shape[2], 0, 0], 'constant', padding_value)
    return {'mel': pad(mel), 'mel_post': pad(mel_post), 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}


# --------------------------------------------------------------------------
# MultiForwardTacotron   (models/multi_forward_tacotron.py)
# --------------------------------------------------------------------------

def _with_speaker(e: torch.Tensor, semb: torch.Tensor) -> torch.Tensor:
    return torch.cat([e, semb[:, None, :].expand(-1, e.shape[1], -1)], dim=2)


@torch.no_grad()
def mft_generate(sd: SD, x: torch.Tensor, speaker_emb: torch.Tensor, alpha: float = 1.0,
                 pitch_function=Identity, energy_function=Identity,
                 pitch_strength: float = 1.0, energy_strength: float = 1.0) -> Dict[str, torch.Tensor]:
    """MultiForwardTacotron.generate/_generate_mel, multi_forward_tacotron.py:243-305.

    Documented deviation: the reference's
    `argmax(pitch_cond_hat.squeeze(), dim=1).unsqueeze(0)` (:252) is only well-formed
    for B=1; here the arg-max is taken over the class axis per (b,t), which is the
    same thing at B=1.  'pitch_cond' is returned as [B,1,T] (== reference at B=1).
    """
    def plain(p):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        return series_predictor(sd, p, _with_speaker(e, speaker_emb))

    def conditional(p, cond, a=1.0):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        c = F.embedding(cond, sd[p + '.pitch_cond_embedding.weight'])
        return series_predictor(sd, p, _with_speaker(torch.cat([e, c], dim=2), speaker_emb), a)

    pc_logits = plain('pitch_cond_pred')                                   # [B,T,3]
    pitch_cond = torch.argmax(pc_logits, dim=-1).long()                     # [B,T]
    dur_hat = apply_fallback(conditional('dur_pred', pitch_cond, alpha).squeeze(-1))
    pitch_hat = pitch_function(conditional('pitch_pred', pitch_cond).transpose(1, 2))
    energy_hat = energy_function(plain('energy_pred').transpose(1, 2))

    e = F.embedding(x, sd['embedding.weight']).transpose(1, 2)
    h = _with_speaker(cbhg(sd, 'prenet', e), speaker_emb)                   # :278-282
    h = h + cond_projection(sd, 'pitch_proj', pitch_hat) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', energy_hat) * energy_strength
    expanded, _, lens, dur_clamped = length_regulate_np(h.numpy(), dur_hat.numpy())
    h = bi_rnn(sd, 'lstm', torch.from_numpy(expanded), 'lstm')
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    post = cbhg(sd, 'postnet', mel)
    mel_post = linear(post, sd['post_proj.weight']).transpose(1, 2)
    return {'mel': mel, 'mel_post': mel_post, 'dur': torch.from_numpy(dur_clamped),
            'pitch': pitch_hat, 'energy': energy_hat, 'pitch_cond': pitch_cond.unsqueeze(1),
            'pitch_cond_logits': pc_logits, 'mel_lens': torch.from_numpy(lens)}


@torch.no_grad()
def mft_forward(sd: SD, batch: Dict[str, torch.Tensor], pitch_strength: float = 1.0, energy_strength: float = 1.0,
                padding_value: float = -11.5129) -> Dict[str, torch.Tensor]:
    """MultiForwardTacotron.forward(batch) in eval mode (multi_forward_tacotron.py:186-241): the dur / pitch predictors
    are conditioned on the TARGET pitch condition batch['pitch_cond'] (:199-200), the mel branch is teacher forced and
    runs the BiLSTM with packed-sequence semantics over `mel_len`."""
    x, semb, cond = batch['x'], batch['speaker_emb'], batch['pitch_cond']

    def plain(p):
        return series_predictor(sd, p, _with_speaker(F.embedding(x, sd[p + '.embedding.weight']), semb))

    def conditional(p):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        c = F.embedding(cond, sd[p + '.pitch_cond_embedding.weight'])
        return series_predictor(sd, p, _with_speaker(torch.cat([e, c], dim=2), semb))

    pitch_cond_hat = plain('pitch_cond_pred').squeeze(-1)                    # [B,T,3]
    dur_hat = conditional('dur_pred').squeeze(-1)
    pitch_hat = conditional('pitch_pred').transpose(1, 2)
    energy_hat = plain('energy_pred').transpose(1, 2)
    h = _with_speaker(cbhg(sd, 'prenet', F.embedding(x, sd['embedding.weight']).transpose(1, 2)), semb)
    h = h + cond_projection(sd, 'pitch_proj', batch['pitch'].unsqueeze(1)) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', batch['energy'].unsqueeze(1)) * energy_strength
    expanded, _, _, _ = length_regulate_np(h.numpy(), batch['dur'].to(torch.float32).numpy().copy())
    h = packed_bi_rnn(sd, 'lstm', torch.from_numpy(expanded), batch['mel_len'], 'lstm', padding_value)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    mel_post = linear(cbhg(sd, 'postnet', mel), sd['post_proj.weight']).transpose(1, 2)
    target = batch['mel'].shape[2]

    def pad(t):
        t = t[:, :, :target]
        return F.pad(t, [0, target - t.shape[2], 0, 0], 'constant', padding_value)
    return {'mel': pad(mel), 'mel_post': pad(mel_post), 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat,
            'pitch_cond': pitch_cond_hat}


# --------------------------------------------------------------------------
# FastPitch   (models/fast_pitch.py, FFT blocks in common_layers.py:127-223)
# --------------------------------------------------------------------------

def layer_norm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + 1e-5) * w + b


def self_attention(sd: SD, p: str, x: torch.Tensor, heads: int,
                   key_pad: Optional[torch.Tensor]) -> torch.Tensor:
    """nn.MultiheadAttention forward (packed in_proj, scale 1/sqrt(d_head), softmax over
    keys, key_padding_mask -> -inf), as used by FFTBlock (common_layers.py:173-175).
    x [B,L,E] (batch-first here; the reference is seq-first, same math)."""
    B, L, E = x.shape
    d = E // heads
    qkv = linear(x, sd[p + '.in_proj_weight'], sd[p + '.in_proj_bias'])
    q, k, v = qkv.chunk(3, dim=-1)
    q = q.view(B, L, heads, d).transpose(1, 2)
    k = k.view(B, L, heads, d).transpose(1, 2)
    v = v.view(B, L, heads, d).transpose(1, 2)
    s = torch.matmul(_QUANT(q) / math.sqrt(d), _QUANT(k).transpose(-1, -2))
    if key_pad is not None:
        s = s.masked_fill(key_pad[:, None, None, :], float('-inf'))
    a = torch.softmax(s, dim=-1)
    o = torch.matmul(_QUANT(a), _QUANT(v)).transpose(1, 2).reshape(B, L, E)
    return linear(o, sd[p + '.out_proj.weight'], sd[p + '.out_proj.bias'])


def fft_block(sd: SD, p: str, x: torch.Tensor, heads: int, key_pad) -> torch.Tensor:
    """FFTBlock.forward (post-norm), common_layers.py:170-185."""
    x = layer_norm(x + self_attention(sd, p + '.self_attn', x, heads, key_pad),
                   sd[p + '.norm1.weight'], sd[p + '.norm1.bias'])
    w1, w2 = sd[p + '.conv1.weight'], sd[p + '.conv2.weight']
    y = F.conv1d(_QUANT(x.transpose(1, 2)), _QUANT(w1), sd[p + '.conv1.bias'], padding=w1.shape[2] // 2)
    y = F.relu(y)
    y = F.conv1d(_QUANT(y), _QUANT(w2), sd[p + '.conv2.bias'], padding=w2.shape[2] // 2)
    return layer_norm(x + y.transpose(1, 2), sd[p + '.norm2.weight'], sd[p + '.norm2.bias'])


def forward_transformer(sd: SD, p: str, x: torch.Tensor, heads: int, key_pad=None) -> torch.Tensor:
    """ForwardTransformer.forward, common_layers.py:211-223.  x [B,L,E]."""
    L = x.shape[1]
    pe = sd[p + '.pos_encoder.pe'][:L, 0, :]                                 # [L,E]
    x = x + sd[p + '.pos_encoder.scale'] * pe[None]                          # :144
    i = 0
    while f'{p}.layers.{i}.norm1.weight' in sd:
        x = fft_block(sd, f'{p}.layers.{i}', x, heads, key_pad)
        i += 1
    return layer_norm(x, sd[p + '.norm.weight'], sd[p + '.norm.bias'])


@torch.no_grad()
def fp_generate(sd: SD, x: torch.Tensor, alpha: float = 1.0,
                pitch_function=Identity, energy_function=Identity,
                pitch_strength: float = 1.0, energy_strength: float = 1.0,
                heads: int = 2) -> Dict[str, torch.Tensor]:
    """FastPitch.generate/_generate_mel, fast_pitch.py:167-221.  Predictors run WITHOUT a
    pad mask (:174-180); the prenet masks keys where x == 0 (:200); the postnet has no
    mask (:215); 'mel' and 'mel_post' are the same tensor (:220)."""
    def pred(p, a=1.0):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        y = forward_transformer(sd, p + '.transformer', e, heads)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias']) / a
    dur_hat = apply_fallback(pred('dur_pred', alpha).squeeze(2))
    pitch_hat = pitch_function(pred('pitch_pred').transpose(1, 2))
    energy_hat = energy_function(pred('energy_pred').transpose(1, 2))
    key_pad = (x == 0)
    h = forward_transformer(sd, 'prenet', F.embedding(x, sd['embedding.weight']), heads, key_pad)
    h = h + cond_projection(sd, 'pitch_proj', pitch_hat) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', energy_hat) * energy_strength
    expanded, _, lens, dur_clamped = length_regulate_np(h.numpy(), dur_hat.numpy())
    h = forward_transformer(sd, 'postnet', torch.from_numpy(expanded), heads, None)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    return {'mel': mel, 'mel_post': mel, 'dur': torch.from_numpy(dur_clamped),
            'pitch': pitch_hat, 'energy': energy_hat, 'mel_lens': torch.from_numpy(lens)}


@torch.no_grad()
def fp_forward(sd: SD, batch: Dict[str, torch.Tensor], pitch_strength: float = 1.0, energy_strength: float = 1.0,
               heads: int = 2, padding_value: float = -11.5129) -> Dict[str, torch.Tensor]:
    """FastPitch.forward(batch) in eval mode (fast_pitch.py:122-165): predictors AND prenet with the token pad mask
    (x == 0), teacher-forced durations / pitch / energy, postnet with the frame mask m >= mel_len[b], outputs cut /
    padded to the target mel length with `padding_value`."""
    x = batch['x']
    key_pad = (x == 0)

    def pred(p):
        y = forward_transformer(sd, p + '.transformer', F.embedding(x, sd[p + '.embedding.weight']), heads, key_pad)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias'])
    dur_hat = pred('dur_pred').squeeze(-1)
    pitch_hat = pred('pitch_pred').transpose(1, 2)
    energy_hat = pred('energy_pred').transpose(1, 2)
    h = forward_transformer(sd, 'prenet', F.embedding(x, sd['embedding.weight']), heads, key_pad)
    h = h + cond_projection(sd, 'pitch_proj', batch['pitch'].unsqueeze(1)) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', batch['energy'].unsqueeze(1)) * energy_strength
    expanded, _, _, _ = length_regulate_np(h.numpy(), batch['dur'].to(torch.float32).numpy().copy())
    M = expanded.shape[1]
    frame_pad = torch.arange(M)[None, :] >= batch['mel_len'][:, None]
    h = forward_transformer(sd, 'postnet', torch.from_numpy(expanded), heads, frame_pad)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    target = batch['mel'].shape[2]
    mel = F.pad(mel[:, :, :target], [0, max(0, target - mel.shape[2]), 0, 0], 'constant', padding_value)
    return {'mel': mel, 'mel_post': mel, 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat}


# --------------------------------------------------------------------------
# MultiFastPitch   (models/multi_fast_pitch.py)
# --------------------------------------------------------------------------

@torch.no_grad()
def mfp_generate(sd: SD, x: torch.Tensor, speaker_emb: torch.Tensor, alpha: float = 1.0,
                 pitch_function=Identity, energy_function=Identity,
                 pitch_strength: float = 1.0, energy_strength: float = 1.0,
                 heads: int = 2) -> Dict[str, torch.Tensor]:
    """MultiFastPitch.generate/_generate_mel, multi_fast_pitch.py:233-305.

    Predictor inputs are concatenations [embedding(x), conditional_embedding(pitch_cond), speaker_emb] (:76-87,
    :43-48) fed to a ForwardTransformer WITHOUT a pad mask (:242-251); the main branch concatenates the speaker
    embedding BEFORE the prenet (:277-283, key mask x == 0) and the postnet runs unmasked (:295).
    Documented deviation (same as MultiForwardTacotron): the reference's
    `argmax(pitch_cond_hat.squeeze(), dim=1).unsqueeze(0)` (:243) is only well-formed for B = 1; here the arg-max is
    per (b,t) over the class axis.  'pitch_cond' is returned as [B,T] (== the reference's [1,T] at B = 1).
    """
    def plain(p, a=1.0):
        e = _with_speaker(F.embedding(x, sd[p + '.embedding.weight']), speaker_emb)
        y = forward_transformer(sd, p + '.transformer', e, heads)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias']) / a

    def conditional(p, cond, a=1.0):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        c = F.embedding(cond, sd[p + '.conditional_embedding.weight'])
        y = forward_transformer(sd, p + '.transformer', _with_speaker(torch.cat([e, c], dim=2), speaker_emb), heads)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias']) / a

    pc_logits = plain('pitch_cond_pred', alpha)                              # [B,T,3]  (:242, divided by alpha)
    pitch_cond = torch.argmax(pc_logits, dim=-1).long()                      # [B,T]
    dur_hat = apply_fallback(conditional('dur_pred', pitch_cond, alpha).squeeze(2))
    pitch_hat = pitch_function(conditional('pitch_pred', pitch_cond).transpose(1, 2))
    energy_hat = energy_function(plain('energy_pred').transpose(1, 2))
    key_pad = (x == 0)
    h = _with_speaker(F.embedding(x, sd['embedding.weight']), speaker_emb)
    h = forward_transformer(sd, 'prenet', h, heads, key_pad)
    h = h + cond_projection(sd, 'pitch_proj', pitch_hat) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', energy_hat) * energy_strength
    expanded, _, lens, dur_clamped = length_regulate_np(h.numpy(), dur_hat.numpy())
    h = forward_transformer(sd, 'postnet', torch.from_numpy(expanded), heads, None)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    return {'mel': mel, 'mel_post': mel, 'dur': torch.from_numpy(dur_clamped), 'pitch': pitch_hat,
            'energy': energy_hat, 'pitch_cond': pitch_cond, 'pitch_cond_logits': pc_logits,
            'mel_lens': torch.from_numpy(lens)}


@torch.no_grad()
def mfp_forward(sd: SD, batch: Dict[str, torch.Tensor], pitch_strength: float = 1.0, energy_strength: float = 1.0,
                heads: int = 2, padding_value: float = -11.5129) -> Dict[str, torch.Tensor]:
    """MultiFastPitch.forward(batch) in eval mode (multi_fast_pitch.py:192-245): every predictor AND the prenet run
    with the token pad mask (x == 0, :205-218), dur / pitch are conditioned on the TARGET pitch condition, the mel
    branch is teacher forced and the postnet runs with the frame mask m >= mel_len[b] (:230-234)."""
    x, semb, cond = batch['x'], batch['speaker_emb'], batch['pitch_cond']
    key_pad = (x == 0)

    def plain(p):
        e = _with_speaker(F.embedding(x, sd[p + '.embedding.weight']), semb)
        y = forward_transformer(sd, p + '.transformer', e, heads, key_pad)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias'])

    def conditional(p):
        e = F.embedding(x, sd[p + '.embedding.weight'])
        c = F.embedding(cond, sd[p + '.conditional_embedding.weight'])
        y = forward_transformer(sd, p + '.transformer', _with_speaker(torch.cat([e, c], dim=2), semb), heads, key_pad)
        return linear(y, sd[p + '.lin.weight'], sd[p + '.lin.bias'])

    dur_hat = conditional('dur_pred').squeeze(-1)
    pitch_hat = conditional('pitch_pred').transpose(1, 2)
    pitch_cond_hat = plain('pitch_cond_pred')                                # [B,T,3]
    energy_hat = plain('energy_pred').transpose(1, 2)
    h = forward_transformer(sd, 'prenet', _with_speaker(F.embedding(x, sd['embedding.weight']), semb), heads, key_pad)
    h = h + cond_projection(sd, 'pitch_proj', batch['pitch'].unsqueeze(1)) * pitch_strength
    h = h + cond_projection(sd, 'energy_proj', batch['energy'].unsqueeze(1)) * energy_strength
    expanded, _, _, _ = length_regulate_np(h.numpy(), batch['dur'].to(torch.float32).numpy().copy())
    M = expanded.shape[1]
    frame_pad = torch.arange(M)[None, :] >= batch['mel_len'][:, None]
    h = forward_transformer(sd, 'postnet', torch.from_numpy(expanded), heads, frame_pad)
    mel = linear(h, sd['lin.weight'], sd['lin.bias']).transpose(1, 2)
    target = batch['mel'].shape[2]
    mel = F.pad(mel[:, :, :target], [0, max(0, target - mel.shape[2]), 0, 0], 'constant', padding_value)
    return {'mel': mel, 'mel_post': mel, 'dur': dur_hat, 'pitch': pitch_hat, 'energy': energy_hat,
            'pitch_cond': pitch_cond_hat}
