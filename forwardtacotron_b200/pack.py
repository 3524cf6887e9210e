"""Weight repacking: torch parameters -> the layouts libftb200.so consumes.

Derived, cached data only: the nn.Parameters stay the source of truth (state_dict keys are
unchanged) and a pack is rebuilt whenever the parameters' versions change.
All GEMM operands are rounded to tf32 (round-to-nearest, ties away: what cvt.rna.tf32.f32
does) here, once, so the tensor core's operand truncation is exact.
"""
from typing import Optional

import torch

BN_EPS = 1e-5


def round_tf32(t: torch.Tensor) -> torch.Tensor:
    """fp32 -> nearest tf32 (10-bit mantissa), ties away from zero; bit-identical to cvt.rna.tf32.f32
    for finite inputs."""
    i = t.detach().contiguous().to(torch.float32).view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


def ceil_to(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def pick_bn(n: int) -> int:
    """N-tile width of the tcgen05 kernel: a multiple of 16, at most 256, chosen to minimise the padded
    width ceil(n/bn)*bn (ties -> the widest tile): 384 -> 192 x 2, 80 -> 80, 768 -> 256 x 3."""
    best = None
    for bn in range(256, 15, -16):
        pad = ceil_to(n, bn)
        if best is None or pad < best[0]:
            best = (pad, bn)
    return best[1]


class PackedGemm:
    """A conv / linear weight as [taps][n_pad][cin] (K-major, tf32-rounded) + epilogue vectors."""

    def __init__(self, w: torch.Tensor, taps: int, tap0: int, n: int, bn: int,
                 bias0: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None,
                 shift: Optional[torch.Tensor] = None, relu: bool = False):
        self.w, self.taps, self.tap0, self.n, self.bn = w, taps, tap0, n, bn
        self.n_pad = w.shape[1]
        self.cin = w.shape[2]
        self.bias0, self.scale, self.shift, self.relu = bias0, scale, shift, relu
        self._w16 = None
        self.w_src = None            # unrounded fp32 copy of `w`; kept only by packs built with split=True
        self._w16s = None
        self._w16f = None

    @property
    def w16(self) -> torch.Tensor:
        """The same packed weight in fp16 (FTB_F_A_F16 problems): tf32-rounded values have an 11-bit significand, so
        the conversion is exact inside fp16's exponent range (|w| < 6e-5 loses relative, not absolute, precision)."""
        if self._w16 is None:
            assert self.cin % 8 == 0, 'fp16 operands need cin % 8 == 0 (16-byte TMA rows)'
            self._w16 = self.w.to(torch.float16).contiguous()
        return self._w16


    @property
    def w16f(self) -> torch.Tensor:
        """The fp16 weight with the taps FOLDED into the reduction (FTB_F_TAPS_IN_K): [n_pad][taps*cin], column
        j*cin + c = tap j, channel c -- the operand of ONE tap over an overlapping-row view of a dense activation."""
        if self._w16f is None:
            self._w16f = self.w16.permute(1, 0, 2).reshape(self.n_pad, self.taps * self.cin).contiguous()
        return self._w16f

    @property
    def w16s(self) -> torch.Tensor:
        """Split-precision weight for a [hi | lo | hi] operand (FTB_F_OUT_SPLIT producers, embed mode 3): per tap
        [n_pad][3*seg] = [W_hi | W_hi | W_lo] with seg = cin rounded up to 8, W_hi = fp16(W), W_lo = fp16(W - W_hi), so
        one kind::f16 GEMM over K = 3*seg accumulates a_hi*W_hi + a_lo*W_hi + a_hi*W_lo."""
        if self._w16s is None:
            assert self.w_src is not None, 'pack with split=True'
            seg = ceil_to(self.cin, 8)
            hi = self.w_src.to(torch.float16)
            lo = (self.w_src - hi.float()).to(torch.float16)
            w = self.w_src.new_zeros((self.taps, self.n_pad, 3 * seg), dtype=torch.float16)
            w[:, :, :self.cin] = hi
            w[:, :, seg:seg + self.cin] = hi
            w[:, :, 2 * seg:2 * seg + self.cin] = lo
            self._w16s = w.contiguous()
        return self._w16s


def _f(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if t is None else t.detach().to(torch.float32).contiguous()


def pack_conv(weight: torch.Tensor, bias: Optional[torch.Tensor] = None, bn: Optional[int] = None,
              scale=None, shift=None, relu=False, split=False) -> PackedGemm:
    """Conv1d weight [Cout, Cin, k] (padding k//2, common_layers.py:50) -> taps x [Cout_pad, Cin]."""
    cout, cin, k = weight.shape
    bn = bn or pick_bn(cout)
    n_pad = ceil_to(cout, bn)
    w = weight.new_zeros((k, n_pad, cin), dtype=torch.float32)
    w[:, :cout, :] = weight.detach().permute(2, 0, 1)
    g = PackedGemm(round_tf32(w), k, -(k // 2), cout, bn, _f(bias), _f(scale), _f(shift), relu)
    if split:
        g.w_src = w
    return g


def pack_bn_conv(conv_w, gamma, beta, mean, var, relu: bool, bn: Optional[int] = None, split=False) -> PackedGemm:
    """BatchNormConv (conv -> ReLU? -> BN eval, common_layers.py:54-57): BN becomes the per-channel
    affine y = v*s + b applied AFTER the ReLU in the epilogue (it cannot be folded into the weights
    when relu=True)."""
    s = gamma.detach().float() / torch.sqrt(var.detach().float() + BN_EPS)
    b = beta.detach().float() - mean.detach().float() * s
    return pack_conv(conv_w, None, bn, s, b, relu, split)


def pack_linear(weight: torch.Tensor, bias: Optional[torch.Tensor] = None, bn: Optional[int] = None,
                relu=False, split=False) -> PackedGemm:
    return pack_conv(weight.detach().unsqueeze(2), bias, bn, None, None, relu, split)


def pack_highway(w1, b1, w2, b2, bn: int = 256) -> PackedGemm:
    """HighwayNetwork (common_layers.py:27-40) as ONE GEMM with N = 2*units: N-tile i holds
    [W1 rows | W2 rows] of units [i*bn/2, (i+1)*bn/2) so both halves of a unit land in one TMEM tile."""
    units, k = w1.shape
    u = bn // 2
    assert units % u == 0, 'highway width must be a multiple of bn/2'
    blocks = []
    for i in range(units // u):
        blocks += [w1.detach()[i * u:(i + 1) * u], w2.detach()[i * u:(i + 1) * u]]
    w = torch.cat(blocks, dim=0).unsqueeze(0).float()
    bias = torch.cat([b1.detach(), b2.detach()]).float().contiguous()
    return PackedGemm(round_tf32(w), 1, 0, 2 * units, bn, bias)


class PackedRnn:
    """Bidirectional GRU/LSTM: input projection as one GEMM over both directions, W_hh resident."""

    def __init__(self, rnn: torch.nn.RNNBase, precise: bool = False):
        """precise: the split-precision form (duration predictors): W_ih packed for a [hi | lo | hi] operand, the
        recurrence run with ftb_rnn_problem.precise."""
        self.precise = precise
        p = dict(rnn.named_parameters())
        self.cell = 1 if isinstance(rnn, torch.nn.LSTM) else 0
        self.hidden = rnn.hidden_size
        w_ih = torch.cat([p['weight_ih_l0'], p['weight_ih_l0_reverse']], dim=0)
        b_ih = torch.cat([p['bias_ih_l0'], p['bias_ih_l0_reverse']], dim=0)
        self.ih = pack_linear(w_ih, b_ih, split=precise)
        # fp32 here: the recurrence kernel rounds W_hh ONCE to fp16 (same 11 significant bits as tf32) when it
        # builds its register-resident MMA fragments
        self.w_hh = torch.stack([p['weight_hh_l0'], p['weight_hh_l0_reverse']]).detach().float().contiguous()
        self.b_hh = torch.stack([p['bias_hh_l0'], p['bias_hh_l0_reverse']]).detach().float().contiguous()
