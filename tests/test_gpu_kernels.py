"""GPU parity tests of the individual sm_100a kernels, called through the C ABI.

Each tensor-core case is checked three ways: tcgen05 kernel vs a float64 torch restatement
of the same (tf32-rounded) operands, and vs the library's own SIMT cross-check kernel.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from forwardtacotron_b200 import engine as eng            # noqa: E402
from forwardtacotron_b200 import pack as pk               # noqa: E402
from forwardtacotron_b200 import runtime as rt            # noqa: E402
from oracle import ft_oracle as O                         # noqa: E402
from tests.helpers import load_golden, max_abs            # noqa: E402

DEV = torch.device('cuda', 0)


def _handle():
    return rt.Handle.get(DEV)


def _fill(prog, name, cols, gen):
    """Random tf32-rounded activation in the valid rows of a padded-flat buffer."""
    g = prog.geo
    buf = prog.buf(name, cols)
    v = pk.round_tf32(torch.randn(g.batch, g.seq_len, cols, generator=gen))
    buf.view(g.batch, g.stride, cols)[:, :g.seq_len] = v.to(DEV)
    return buf, v.double()


def _conv_ref(x, w, k):
    """x [B,L,Cin] float64, w [Cout,Cin,k] -> [B,L,Cout] with padding k//2 trimmed to L."""
    y = F.conv1d(x.transpose(1, 2), w.double(), padding=k // 2)[:, :, :x.shape[1]]
    return y.transpose(1, 2)


@pytest.mark.parametrize('B,L,cin,cout,k,relu', [
    (2, 50, 64, 256, 5, True),        # predictor first conv
    (3, 130, 256, 256, 1, True),      # k=1, crosses an M-tile boundary
    (2, 70, 80, 256, 4, True),        # even k (asymmetric trim), Cin=80 (K tail zero-filled by TMA)
    (1, 300, 256, 80, 3, False),      # N=80, no ReLU
    (2, 40, 388, 256, 5, True),       # multi-speaker predictor width
    (2, 40, 392, 1176, 1, False),     # MultiFastPitch in_proj: K tail AND N not a multiple of 16
    (2, 40, 392, 392, 1, False),      # MultiFastPitch out_proj / conv2 width
    (1, 70, 392, 128, 9, True),       # MultiFastPitch conv1 (k = 9)
])
def test_conv_bn(B, L, cin, cout, k, relu):
    gen = torch.Generator().manual_seed(B * 1000 + L + k)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a, a_ref = _fill(prog, 'a', cin, gen)
    w = torch.randn(cout, cin, k, generator=gen) / (cin * k) ** 0.5
    gamma, beta = torch.rand(cout, generator=gen) + 0.5, torch.randn(cout, generator=gen) * 0.1
    mean, var = torch.randn(cout, generator=gen) * 0.1, torch.rand(cout, generator=gen) + 0.5
    packed = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), relu)
    out = prog.buf('out', cout)
    prog.gemm(a, packed, out, round_out=False)
    prog.flush()
    y = _conv_ref(a_ref, pk.round_tf32(w), k)
    if relu:
        y = y.clamp_min(0)
    s = (gamma / torch.sqrt(var + 1e-5)).double()
    ref = y * s + (beta.double() - mean.double() * s)
    for impl in (1, 0):
        out.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        got = prog.view('out').cpu().double()
        assert max_abs(got, ref) < 2e-4, f'impl={impl}'
    prog.h.impl = 0
    # gap rows stay zero
    g = prog.geo
    assert float(out.view(g.batch, g.stride, cout)[:, g.seq_len:].abs().max()) == 0.0


@pytest.mark.parametrize('B,L,cin,cout,k,relu,pool,in16,out16', [
    (2, 130, 256, 256, 3, True, False, True, False),     # fp16 operands, fp32 result (CBHG proj2)
    (2, 130, 2048, 256, 3, True, False, True, True),      # fp16 operands and result (CBHG proj1, K = 3 x 2048)
    (3, 70, 80, 256, 4, True, True, False, True),         # tf32 operands, pooled fp16 result (postnet bank conv)
    (2, 200, 256, 128, 16, True, True, True, True),       # fp16 both sides, pooled, N = 128 < one chunk group
    (1, 50, 72, 48, 5, False, False, True, True),         # K tail (72 of 128 per k-block), half chunk (bn = 48)
])
def test_conv_fp16_operands(B, L, cin, cout, k, relu, pool, in16, out16):
    """FTB_F_A_F16 / FTB_F_OUT_F16: fp16 GEMM-to-GEMM hand-off buffers (kind::f16 MMAs, fp16 row-segment stores)."""
    gen = torch.Generator().manual_seed(B * 100 + L + k)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    v = torch.randn(B, L, cin, generator=gen).to(torch.float16 if in16 else torch.float32)
    if not in16:
        v = pk.round_tf32(v)
    a = prog.buf('a', cin, torch.float16 if in16 else torch.float32)
    a.view(g.batch, g.stride, cin)[:, :L] = v.to(DEV)
    w = torch.randn(cout, cin, k, generator=gen) / (cin * k) ** 0.5
    gamma, beta = torch.rand(cout, generator=gen) + 0.5, torch.randn(cout, generator=gen) * 0.1
    mean, var = torch.randn(cout, generator=gen) * 0.1, torch.rand(cout, generator=gen) + 0.5
    packed = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), relu)
    out = prog.buf('out', cout + 8, torch.float16 if out16 else torch.float32)     # written at column 8
    prog.gemm(a, packed, out, out_col0=8, epilogue=rt.EPI_MAXPOOL if pool else rt.EPI_AFFINE, round_out=False)
    prog.flush()
    w_used = (packed.w16 if in16 else packed.w).cpu().double()[:, :cout, :].permute(1, 2, 0)     # [cout, cin, k]
    y = _conv_ref(v.double(), w_used, k)
    if relu:
        y = y.clamp_min(0)
    sc = (gamma / torch.sqrt(var + 1e-5)).double()
    ref = y * sc + (beta.double() - mean.double() * sc)
    if pool:
        ref = F.max_pool1d(ref.transpose(1, 2), kernel_size=2, stride=1, padding=1)[:, :, :L].transpose(1, 2)
    tol = 2e-4 + (float(ref.abs().max()) * 2.0 ** -11 if out16 else 0.0)
    for impl in (1, 0):
        out.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        got = out.view(g.batch, g.stride, cout + 8)[:, :L, 8:].cpu().double()
        assert max_abs(got, ref) < tol, f'impl={impl}'
        assert float(out[:, :8].abs().max()) == 0.0 and float(out.view(g.batch, g.stride, -1)[:, L:].abs().max()) == 0.0
    prog.h.impl = 0


@pytest.mark.parametrize('B,L,cin,cout,k,pool', [
    (3, 70, 80, 256, 8, True),          # the postnet bank's widest conv: 10 k-blocks instead of 16, pooled (halo row -1)
    (2, 130, 80, 256, 5, True),         # odd k, crosses an M-tile boundary
    (1, 300, 80, 256, 2, False),        # even k (asymmetric trim): 3 k-blocks instead of 4
    (2, 50, 72, 48, 3, False),          # K = 216: tail of the last k-block zero-filled, half chunk (bn = 48)
])
def test_conv_taps_folded_into_k(B, L, cin, cout, k, pool):
    """FTB_F_TAPS_IN_K: ONE tap over an overlapping-row tensor map of a dense fp16 activation with zero lead rows
    (common_layers.py:97-99 over the 80-channel mel).  Checked against float64 torch, the SIMT cross-check kernel and --
    bit for bit -- the unfolded problem on the same operands (same products, same k order inside a tap... the sums differ
    only by fp32 association, so that comparison is to 1e-5)."""
    gen = torch.Generator().manual_seed(B * 10 + L + k)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    v = torch.randn(B, L, cin, generator=gen).to(torch.float16)
    a = prog.buf('a', cin, torch.float16, lead=rt.FOLD_LEAD_ROWS)
    a.view(g.batch, g.stride, cin)[:, :L] = v.to(DEV)
    a2 = prog.buf('a_nolead', cin, torch.float16)                      # same values, no lead rows: stays unfolded
    a2.copy_(a)
    w = torch.randn(cout, cin, k, generator=gen) / (cin * k) ** 0.5
    gamma, beta = torch.rand(cout, generator=gen) + 0.5, torch.randn(cout, generator=gen) * 0.1
    mean, var = torch.randn(cout, generator=gen) * 0.1, torch.rand(cout, generator=gen) + 0.5
    packed = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), True)
    out = prog.buf('out', cout, torch.float16)
    out2 = prog.buf('out_unfolded', cout, torch.float16)
    epi = rt.EPI_MAXPOOL if pool else rt.EPI_AFFINE
    prog.gemm(a, packed, out, epilogue=epi, round_out=False)
    assert prog._pending[-1].flags & rt.F_TAPS_IN_K
    prog.flush()
    prog.gemm(a2, packed, out2, epilogue=epi, round_out=False)
    assert not (prog._pending[-1].flags & rt.F_TAPS_IN_K)
    prog.flush()
    w_used = packed.w16.cpu().double()[:, :cout, :].permute(1, 2, 0)
    y = _conv_ref(v.double(), w_used, k).clamp_min(0)
    sc = (gamma / torch.sqrt(var + 1e-5)).double()
    ref = y * sc + (beta.double() - mean.double() * sc)
    if pool:
        ref = F.max_pool1d(ref.transpose(1, 2), kernel_size=2, stride=1, padding=1)[:, :, :L].transpose(1, 2)
    tol = 2e-4 + float(ref.abs().max()) * 2.0 ** -11
    for impl in (1, 0):
        out.zero_(); out2.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        got = prog.view('out').cpu().double()
        assert max_abs(got, ref) < tol, f'impl={impl}'
        assert max_abs(got, prog.view('out_unfolded').cpu().double()) < 1e-5 + float(ref.abs().max()) * 2.0 ** -10, f'impl={impl}'
        assert float(out.view(g.batch, g.stride, cout)[:, L:].abs().max()) == 0.0
    prog.h.impl = 0
    # another length inside the same capacity: the folded view needs no re-encoding
    L2 = L - 7
    prog.set_seq_len(L2)
    a.zero_()
    a[:g.rows].view(g.batch, g.stride, cin)[:, :L2] = v[:, :L2].to(DEV)
    out.zero_()
    prog.run()
    torch.cuda.synchronize()
    y2 = _conv_ref(v[:, :L2].double(), w_used, k).clamp_min(0)
    ref2 = y2 * sc + (beta.double() - mean.double() * sc)
    if pool:
        ref2 = F.max_pool1d(ref2.transpose(1, 2), kernel_size=2, stride=1, padding=1)[:, :, :L2].transpose(1, 2)
    assert max_abs(prog.view('out').cpu().double(), ref2) < tol


def test_bank_maxpool_matches_reference_semantics():
    """conv bank (k=1..6) + MaxPool1d(2,1,1)[:L] fused epilogue vs torch ops (common_layers.py:97-105)."""
    B, L, cin, ch, K = 2, 200, 64, 64, 6
    gen = torch.Generator().manual_seed(5)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a, a_ref = _fill(prog, 'a', cin, gen)
    bank = prog.buf('bank', K * ch)
    refs = []
    for k in range(1, K + 1):
        w = torch.randn(ch, cin, k, generator=gen) / (cin * k) ** 0.5
        gamma, beta = torch.rand(ch, generator=gen) + 0.5, torch.randn(ch, generator=gen) * 0.3
        mean, var = torch.randn(ch, generator=gen) * 0.1, torch.rand(ch, generator=gen) + 0.5
        packed = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), True)
        prog.gemm(a, packed, bank, out_col0=(k - 1) * ch, epilogue=rt.EPI_MAXPOOL, round_out=False)
        y = _conv_ref(a_ref, pk.round_tf32(w), k).clamp_min(0)
        s = (gamma / torch.sqrt(var + 1e-5)).double()
        refs.append(y * s + (beta.double() - mean.double() * s))
    prog.flush()
    ref = torch.cat(refs, dim=2).transpose(1, 2)                     # [B, K*ch, L]
    ref = F.max_pool1d(ref, kernel_size=2, stride=1, padding=1)[:, :, :L].transpose(1, 2)
    for impl in (1, 0):
        bank.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        assert max_abs(prog.view('bank').cpu().double(), ref) < 2e-4, f'impl={impl}'
    prog.h.impl = 0


def test_highway_and_transposed_linear():
    B, L, ch = 2, 150, 256
    gen = torch.Generator().manual_seed(9)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a, a_ref = _fill(prog, 'a', ch, gen)
    w1, w2 = torch.randn(ch, ch, generator=gen) / 16, torch.randn(ch, ch, generator=gen) / 16
    b1, b2 = torch.randn(ch, generator=gen) * 0.1, torch.randn(ch, generator=gen) * 0.1
    hw = pk.pack_highway(w1.to(DEV), b1.to(DEV), w2.to(DEV), b2.to(DEV))
    out = prog.buf('hw', ch)
    prog.gemm(a, hw, out, epilogue=rt.EPI_HIGHWAY, residual=a, round_out=False)
    wl, bl = torch.randn(80, ch, generator=gen) / 16, torch.randn(80, generator=gen)
    lin = pk.pack_linear(wl.to(DEV), bl.to(DEV))
    mel_cl = prog.buf('mel_cl', 80)
    mel = prog.dense('mel', (B, 80, L))
    prog.gemm(a, lin, mel_cl, out_t=mel, round_out=True)
    prog.flush()
    x1 = a_ref @ pk.round_tf32(w1).double().T + b1.double()
    x2 = a_ref @ pk.round_tf32(w2).double().T + b2.double()
    g = torch.sigmoid(x2)
    ref_hw = g * x1.clamp_min(0) + (1 - g) * a_ref
    ref_mel = a_ref @ pk.round_tf32(wl).double().T + bl.double()
    for impl in (1, 0):
        out.zero_(); mel.zero_(); mel_cl.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        assert max_abs(prog.view('hw').cpu().double(), ref_hw) < 2e-4, f'impl={impl}'
        assert max_abs(mel.cpu().double(), ref_mel.transpose(1, 2)) < 2e-4, f'impl={impl}'
        assert max_abs(prog.view('mel_cl').cpu().double(), ref_mel) < 3e-3      # tf32-rounded copy
    prog.h.impl = 0


def test_gemm_cta_pair_mode_agrees(monkeypatch):
    """FTB_GEMM_PAIR=1 (cta_group::2: M = 256 MMAs issued by the even CTA of a 2-CTA cluster, each CTA staging half of
    the W tile, six-stage ring) against the single-CTA kernel on one launch mixing a pooled fp16 bank convolution, a
    k=3 projection with residual, a highway layer and a transposed linear layer; odd m-tile counts included."""
    B, L, ch = 3, 171, 128                      # 3 * stride rows -> an odd number of m-tiles for the pooled problem
    got = {}
    for mode in ('0', '1'):
        monkeypatch.setenv('FTB_GEMM_PAIR', mode)
        gen = torch.Generator().manual_seed(23)
        prog = eng.Program(_handle(), eng.Geometry(B, L))
        g = prog.geo
        v = torch.randn(B, L, ch, generator=gen)
        a16, a32 = prog.buf('a16', ch, torch.float16), prog.buf('a32', ch)
        a16.view(B, g.stride, ch)[:, :L] = v.to(DEV).half()
        a32.view(B, g.stride, ch)[:, :L] = v.to(DEV)
        w = torch.randn(256, ch, 5, generator=gen) / (ch * 5) ** 0.5
        gamma, beta = torch.rand(256, generator=gen) + 0.5, torch.randn(256, generator=gen) * 0.1
        mean, var = torch.randn(256, generator=gen) * 0.1, torch.rand(256, generator=gen) + 0.5
        bank = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), True)
        o_bank = prog.buf('bank', 256, torch.float16)
        prog.gemm(a16, bank, o_bank, epilogue=rt.EPI_MAXPOOL)
        w3 = torch.randn(ch, ch, 3, generator=gen) / (ch * 3) ** 0.5
        proj = pk.pack_bn_conv(w3.to(DEV), gamma[:ch].to(DEV), beta[:ch].to(DEV), mean[:ch].to(DEV), var[:ch].to(DEV), False)
        o_proj = prog.buf('proj', ch)
        prog.gemm(a16, proj, o_proj, residual=a32)
        w1, w2 = torch.randn(ch, ch, generator=gen) / 11, torch.randn(ch, ch, generator=gen) / 11
        hw = pk.pack_highway(w1.to(DEV), torch.zeros(ch, device=DEV), w2.to(DEV), torch.ones(ch, device=DEV))
        o_hw, o_hwf = prog.buf('hw', ch, torch.float16), prog.buf('hw_full', ch)
        prog.gemm(a16, hw, o_hw, epilogue=rt.EPI_HIGHWAY, residual=a32, out_full=o_hwf)
        lin = pk.pack_linear((torch.randn(80, ch, generator=gen) / 11).to(DEV), torch.randn(80, generator=gen).to(DEV))
        mel = prog.dense('mel', (B, 80, L))
        prog.gemm(a32, lin, None, out_t=mel)
        prog.flush()
        prog.run()
        torch.cuda.synchronize()
        got[mode] = [t.float().cpu().clone() for t in (o_bank, o_proj, o_hw, o_hwf, mel)]
        prog.h.impl = 1                                      # SIMT cross-check of the same plan
        ref = []
        for t in (o_bank, o_proj, o_hw, o_hwf, mel):
            t.zero_()
        prog.run()
        torch.cuda.synchronize()
        prog.h.impl = 0
        for a, t in zip(got[mode], (o_bank, o_proj, o_hw, o_hwf, mel)):
            # fp16 outputs: the two kernels may round a value to neighbouring halves (one ulp = 2^-10 relative)
            assert max_abs(a, t.float().cpu()) < 5e-4 + 2.0 ** -10 * float(a.abs().max()), mode
    for a, b in zip(got['0'], got['1']):
        assert max_abs(a, b) < 1e-5
        assert float(a.abs().max()) > 0.1


def test_fp16_operand_copy_next_to_twin_residual_transposed():
    """FTB_F_OUT_F16 with out_full / residual / highway / out_t: the small-K chain proj2 -> pre_highway -> highways
    -> gru_ih and `lin` hand fp16 operand copies on while their fp32 twins feed the residual terms."""
    B, L, ch = 2, 150, 128
    gen = torch.Generator().manual_seed(19)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    v = torch.randn(B, L, ch, generator=gen)
    a16 = prog.buf('a16', ch, torch.float16)
    a32 = prog.buf('a32', ch)
    a16.view(B, g.stride, ch)[:, :L] = v.to(DEV).half()
    a32.view(B, g.stride, ch)[:, :L] = v.to(DEV)
    a_ref = v.half().double()
    w1, w2 = torch.randn(ch, ch, generator=gen) / 11, torch.randn(ch, ch, generator=gen) / 11
    b1, b2 = torch.randn(ch, generator=gen) * 0.1, torch.randn(ch, generator=gen) * 0.1
    hw = pk.pack_highway(w1.to(DEV), b1.to(DEV), w2.to(DEV), b2.to(DEV))
    hw_o, hw_f = prog.buf('hw', ch, torch.float16), prog.buf('hw_full', ch)
    prog.gemm(a16, hw, hw_o, epilogue=rt.EPI_HIGHWAY, residual=a32, out_full=hw_f)
    wl, bl = torch.randn(80, ch, generator=gen) / 11, torch.randn(80, generator=gen)
    lin = pk.pack_linear(wl.to(DEV), bl.to(DEV))
    mel_cl, mel_f = prog.buf('mel_cl', 80, torch.float16), prog.buf('mel_cl_full', 80)
    mel = prog.dense('mel', (B, 80, L))
    prog.gemm(a16, lin, mel_cl, out_t=mel, out_full=mel_f)
    wp = torch.randn(ch, ch, generator=gen) / 11
    proj = pk.pack_linear(wp.to(DEV), None)
    p_res = prog.buf('p_res', ch, torch.float16)                 # residual only
    p_twin, p_twin_f = prog.buf('p_twin', ch, torch.float16), prog.buf('p_twin_full', ch)
    prog.gemm(a16, proj, p_res, residual=a32)
    prog.gemm(a16, proj, p_twin, out_full=p_twin_f)
    prog.flush()
    r16 = lambda w: w.half().double()        # noqa: E731  (the packed fp16 weight copy of a tf32-rounded weight)
    x1 = a_ref @ r16(pk.round_tf32(w1)).T + b1.double()
    x2 = a_ref @ r16(pk.round_tf32(w2)).T + b2.double()
    gate = torch.sigmoid(x2)
    ref_hw = gate * x1.clamp_min(0) + (1 - gate) * v.double()
    ref_mel = a_ref @ r16(pk.round_tf32(wl)).T + bl.double()
    ref_p = a_ref @ r16(pk.round_tf32(wp)).T
    for impl in (1, 0):
        for b in (hw_o, hw_f, mel_cl, mel_f, mel, p_res, p_twin, p_twin_f):
            b.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        view = lambda n: prog.view(n).cpu().double()             # noqa: E731
        assert max_abs(view('hw_full'), ref_hw) < 2e-4, impl
        assert torch.equal(prog.view('hw'), prog.view('hw_full').half()), impl
        assert max_abs(view('mel_cl_full'), ref_mel) < 2e-4, impl
        assert max_abs(mel.cpu().double(), ref_mel.transpose(1, 2)) < 2e-4, impl
        assert torch.equal(prog.view('mel_cl'), prog.view('mel_cl_full').half()), impl
        assert max_abs(view('p_twin_full'), ref_p) < 2e-4, impl
        assert torch.equal(prog.view('p_twin'), prog.view('p_twin_full').half()), impl
        assert max_abs(view('p_res'), ref_p + v.double()) < 2e-4 + 2.0 ** -11 * float((ref_p + v.double()).abs().max()), impl
        for n in ('hw', 'mel_cl', 'p_res', 'p_twin'):             # pad rows untouched
            assert float(prog.bufs[n].view(B, g.stride, -1)[:, L:].float().abs().max()) == 0.0
    prog.h.impl = 0


@pytest.mark.parametrize('cell,H,B,L', [('gru', 64, 3, 37), ('gru', 128, 5, 20), ('gru', 256, 33, 25),
                                        ('lstm', 512, 2, 30), ('lstm', 512, 40, 12)])
def test_recurrence(cell, H, B, L):
    gen = torch.Generator().manual_seed(H + B)
    In = 96
    rnn = (torch.nn.GRU if cell == 'gru' else torch.nn.LSTM)(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    packed = pk.PackedRnn(rnn.to(DEV))
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a, a_ref = _fill(prog, 'x', In, gen)
    xp = prog.buf('xp', packed.ih.n)
    out = prog.buf('h', 2 * H)
    prog.gemm(a, packed.ih, xp, round_out=False)
    prog.rnn([(packed, xp, out, 0)])
    prog.run()
    torch.cuda.synchronize()
    sd = {'r.' + k: v.detach().cpu() for k, v in rnn.state_dict().items()}
    O.set_operand_rounding(O.round_tf32)
    try:
        with torch.no_grad():
            ref = O.bi_rnn(sd, 'r', a_ref.float(), cell)
    finally:
        O.set_operand_rounding(None)
    got = prog.view('h').cpu()
    assert max_abs(got, ref) < 2e-3
    # second run of the same plan gives the same answer (barrier counters are re-armed)
    out.zero_()
    prog.run()
    torch.cuda.synchronize()
    assert torch.equal(prog.view('h').cpu(), got)


@pytest.mark.parametrize('cell,H,B,L', [('gru', 64, 17, 21), ('gru', 256, 20, 19), ('lstm', 512, 33, 9)])
def test_recurrence_handoff_protocols_agree(cell, H, B, L, monkeypatch):
    """The two hand-off protocols of K3/K4 (peer shared memory inside a thread-block cluster / fp16 exchange tensor in
    L2 with in-band sentinels) run the same arithmetic in the same order: their outputs are bit-identical, and the plan
    reports which one it took (cluster sizes 2, 8 and the non-portable 16 are all exercised here)."""
    gen = torch.Generator().manual_seed(7 * H + B)
    In = 64
    rnn = (torch.nn.GRU if cell == 'gru' else torch.nn.LSTM)(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    packed = pk.PackedRnn(rnn.to(DEV))
    got, gotx = {}, {}
    for mode in ('cluster', 'l2'):
        monkeypatch.setenv('FTB_RNN_HANDOFF', mode)
        prog = eng.Program(_handle(), eng.Geometry(B, L))
        g2 = torch.Generator().manual_seed(B)
        a, _ = _fill(prog, 'x', In, g2)
        xp = prog.buf('xp', packed.ih.n)
        out = prog.buf('h', 2 * H)
        out16 = prog.buf('h16', 2 * H, torch.float16)        # out_f16: the same values stored as fp16 (h_t is fp16-exact)
        prog.gemm(a, packed.ih, xp, round_out=False)
        prog.rnn([(packed, xp, out, 0)])
        assert prog.info[-1]['handoff'] == mode
        prog.rnn([(packed, xp, out16, 0)])
        xp16 = prog.buf('xp16', packed.ih.n, torch.float16)  # x_f16: the x-projection as an fp16 GEMM output
        outx = prog.buf('hx16', 2 * H)
        prog.gemm(a, packed.ih, xp16, round_out=False)
        prog.rnn([(packed, xp16, outx, 0)])
        for _ in range(2):                                   # second run: barriers / sentinels re-armed
            prog.run()
        torch.cuda.synchronize()
        got[mode] = prog.view('h').cpu().clone()
        assert bool(torch.isfinite(got[mode]).all())
        assert torch.equal(prog.view('h16').float().cpu(), got[mode])
        assert float(out16.view(B, prog.geo.stride, -1)[:, L:].float().abs().max()) == 0.0
        gotx[mode] = prog.view('hx16').cpu().clone()
        assert max_abs(gotx[mode], got[mode]) < 5e-3         # one more rounding of the gate pre-activations to 11 bits
        assert max_abs(gotx[mode], got[mode]) > 0.0
    assert torch.equal(got['cluster'], got['l2'])
    assert torch.equal(gotx['cluster'], gotx['l2'])


@pytest.mark.parametrize('cell,H,B,L', [('gru', 128, 5, 33), ('lstm', 512, 19, 27)])
def test_recurrence_packed_sequence_semantics(cell, H, B, L):
    """`lens` != null: pack_padded_sequence -> RNN -> pad_packed_sequence(padding_value) of forward_tacotron.py:147-152
    (each sentence runs over its own length, the backward direction starts at len-1, the tail holds the pad value),
    against torch's packed-sequence RNN in fp32."""
    from torch.nn.utils.rnn import pack_padded_sequence, pad_packed_sequence
    gen = torch.Generator().manual_seed(H + B + 1)
    In, pad = 64, -11.5129
    rnn = (torch.nn.GRU if cell == 'gru' else torch.nn.LSTM)(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    lens = torch.randint(1, L + 1, (B,), generator=gen)
    lens[0] = L                                          # pad_packed_sequence pads to the longest sentence
    a_cpu = pk.round_tf32(torch.randn(B, L, In, generator=gen))
    with torch.no_grad():
        y, _ = rnn(pack_padded_sequence(a_cpu, lens, batch_first=True, enforce_sorted=False))
        ref, _ = pad_packed_sequence(y, batch_first=True, padding_value=pad)
    packed = pk.PackedRnn(rnn.to(DEV))
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a = prog.buf('x', In)
    a.view(B, prog.geo.stride, In)[:, :L] = a_cpu.to(DEV)
    xp = prog.buf('xp', packed.ih.n)
    out = prog.buf('h', 2 * H)
    lens_d = lens.to(torch.int32).to(DEV)
    prog.gemm(a, packed.ih, xp, round_out=False)
    prog.rnn([(packed, xp, out, 0)], lens=lens_d, pad_value=pad)
    prog.run()
    torch.cuda.synchronize()
    got = prog.view('h').cpu()
    assert got.shape == ref.shape
    out16 = prog.buf('h16', 2 * H, torch.float16)
    prog.rnn([(packed, xp, out16, 0)], lens=lens_d, pad_value=pad)
    prog.run()
    torch.cuda.synchronize()
    mask = torch.arange(L)[None, :] >= lens[:, None]
    got16 = prog.view('h16').float().cpu()
    assert torch.equal(got16[~mask], got[~mask]) and bool((got16[mask] == torch.tensor(pad).half().float()).all())
    assert bool((got[mask] == torch.tensor(pad)).all())                      # the tail is exactly the pad value
    assert max_abs(got[~mask], ref[~mask]) < 3e-3                            # tf32 x-projection + fp16 recurrent operands


@pytest.mark.parametrize('tag', ['all2', 'ties', 'zero_row'])
def test_length_regulator_bit_exact_golden(tag):
    g = load_golden(f'lr_{tag}')
    h = _handle()
    x = torch.from_numpy(g['x']).to(DEV)
    B, T, C = x.shape
    dur = torch.from_numpy(g['dur']).to(DEV).contiguous()
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(dur, B, T, False, cum, lens, meta)
    M = int(meta.cpu()[0])
    assert M == g['out'].shape[1]
    assert np.array_equal(dur.cpu().numpy().view(np.uint32), g['dur_after'].view(np.uint32))
    out = torch.full((B, M, C), 7.0, device=DEV)
    idx = torch.empty((B, M), dtype=torch.int32, device=DEV)
    a = rt.LrGatherArgs()
    a.x, a.x_stride, a.x_ld, a.c1, a.c2 = x.data_ptr(), T, C, C, 0
    a.cum, a.lens, a.batch, a.T, a.M = cum.data_ptr(), lens.data_ptr(), B, T, M
    a.out, a.out_stride, a.out_ld, a.idx, a.round_tf32 = out.data_ptr(), M, C, idx.data_ptr(), 0
    h.lr_gather(a)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), g['out'].view(np.uint32))
    _, idx_ref, lens_ref, _ = O.length_regulate_np(g['x'], g['dur'])
    assert np.array_equal(idx.cpu().numpy(), idx_ref)
    assert np.array_equal(lens.cpu().numpy(), lens_ref)
    # mode 2: the fp16 operand copy (what the W_ih GEMM of the BiLSTM reads) is the same gather, rounded to nearest
    if C % 8 == 0:
        out16 = torch.full((B, M, C), 7.0, device=DEV, dtype=torch.float16)
        a.out, a.round_tf32, a.idx = out16.data_ptr(), 2, None
        h.lr_gather(a)
        torch.cuda.synchronize()
        assert torch.equal(out16.cpu(), torch.from_numpy(g['out']).to(torch.float16))


def test_length_regulator_full_size_properties():
    """Config-2 size (64 x 200 x 512): bit-exact vs the numpy oracle + fallback flag."""
    B, T, C = 64, 200, 512
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(B, T, C, generator=gen)
    dur = torch.rand(B, T, generator=gen) * 5 - 0.5
    h = _handle()
    xd, dd = x.to(DEV), dur.to(DEV).contiguous()
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(dd, B, T, True, cum, lens, meta)
    m = meta.cpu()
    ref, idx_ref, lens_ref, _ = O.length_regulate_np(x.numpy(), dur.numpy())
    M = int(m[0])
    assert M == ref.shape[1] and int(m[1]) == 0 and int(m[2]) == int(lens_ref.sum())
    out = torch.empty((B, M, C), device=DEV)
    a = rt.LrGatherArgs()
    a.x, a.x_stride, a.x_ld, a.c1, a.c2 = xd.data_ptr(), T, C, C, 0
    a.cum, a.lens, a.batch, a.T, a.M = cum.data_ptr(), lens.data_ptr(), B, T, M
    a.out, a.out_stride, a.out_ld, a.idx, a.round_tf32 = out.data_ptr(), M, C, None, 0
    h.lr_gather(a)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    # fallback: all durations truncate to <= 0  ->  every duration becomes exactly 2
    d0 = (torch.rand(B, T, generator=gen) * 0.9 - 0.3).to(DEV).contiguous()
    h.lr_scan(d0, B, T, True, cum, lens, meta)
    m = meta.cpu()
    assert int(m[1]) == 1 and int(m[0]) == 2 * T and bool((d0 == 2.0).all())


@pytest.mark.parametrize('D,heads,B,L,masked', [(64, 2, 3, 70, True), (128, 2, 2, 130, False), (192, 2, 2, 50, True),
                                                (196, 2, 2, 77, True), (256, 2, 1, 100, False)])
def test_mha_matches_torch(D, heads, B, L, masked):
    """K8 through the C ABI vs torch scaled-dot-product attention in float64 (head dims of FastPitch and
    MultiFastPitch; key-padding mask = token id 0, common_layers.py:232-233)."""
    gen = torch.Generator().manual_seed(D + L)
    E = D * heads
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    qkv, ref = _fill(prog, 'qkv', 3 * E, gen)
    tok = torch.ones(B, L, dtype=torch.int64)
    if masked:
        for b in range(B):
            tok[b, L - 3 - 5 * b:] = 0
    tok_d = tok.to(DEV)
    out = prog.buf('att', E)
    prog.h.mha(qkv, E, heads, tok_d if masked else None, B, L, prog.geo.stride, out, round_tf32=False)
    torch.cuda.synchronize()
    q, k, v = [t.view(B, L, heads, D).transpose(1, 2) for t in ref.chunk(3, dim=-1)]
    s = q @ k.transpose(-1, -2) / D ** 0.5
    if masked:
        s = s.masked_fill((tok == 0)[:, None, None, :], float('-inf'))
    want = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, L, E)
    assert max_abs(prog.view('att').cpu().double(), want) < 2e-5


@pytest.mark.parametrize('E,chained,with_res', [
    (128, False, False),     # FastPitch predictors: one float4 per lane
    (256, True, False),      # FastPitch pre / postnet, last layer chained with the transformer's final norm
    (384, False, True),      # three float4 per lane, residual operand
    (392, True, True),       # MultiFastPitch width: tail of the last chunk predicated
    (1000, False, False),    # eight chunks per lane
    (102, False, True),      # e % 4 != 0: the scalar kernel
])
def test_add_layernorm_all_output_forms(E, chained, with_res):
    """K7 through the C ABI vs float64 torch LayerNorm (common_layers.py:177,184,209): fp32 as computed, tf32-rounded, the
    fp16 operand copy (mode 2) and the split [hi | lo | hi] form (mode 3), with the unrounded twin next to each."""
    gen = torch.Generator().manual_seed(E)
    B, L = 3, 45
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    x, x_ref = _fill(prog, 'x', E, gen)
    res, res_ref = _fill(prog, 'res', E, gen) if with_res else (None, None)
    g1, b1 = (torch.rand(E, generator=gen) + 0.5).to(DEV), (torch.randn(E, generator=gen) * 0.1).to(DEV)
    g2, b2 = ((torch.rand(E, generator=gen) + 0.5).to(DEV), (torch.randn(E, generator=gen) * 0.1).to(DEV)) if chained else (None, None)
    v = x_ref + (res_ref if with_res else 0.0)
    want = F.layer_norm(v, (E,), g1.cpu().double(), b1.cpu().double(), 1e-5)
    if chained:
        want = F.layer_norm(want, (E,), g2.cpu().double(), b2.cpu().double(), 1e-5)
    h = _handle()
    modes = [0, 1] + ([2, 3] if E % 8 == 0 else [])
    for mode in modes:
        cols, dt = {0: (E, torch.float32), 1: (E, torch.float32), 2: (E, torch.float16), 3: (3 * E, torch.float16)}[mode]
        out = prog.buf(f'out{mode}', cols, dt)
        full = prog.buf(f'full{mode}', E)
        h.add_layernorm(x, res, E, E, g1, b1, g2, b2, B, L, g.stride, out, round_tf32=mode, out_full=full)
        torch.cuda.synchronize()
        assert max_abs(prog.view(f'full{mode}').cpu().double(), want) < 2e-5, mode
        o = prog.view(f'out{mode}').cpu()
        f32 = prog.view(f'full{mode}').cpu()
        if mode == 0:
            assert torch.equal(o, f32)
        elif mode == 1:
            assert torch.equal(o, pk.round_tf32(f32))
        elif mode == 2:
            assert torch.equal(o, f32.to(torch.float16))
        else:
            hi = f32.to(torch.float16)
            assert torch.equal(o[..., :E], hi) and torch.equal(o[..., 2 * E:], hi)
            assert torch.equal(o[..., E:2 * E], (f32 - hi.float()).to(torch.float16))
        assert float(out.view(g.batch, g.stride, -1)[:, L:].abs().max()) == 0.0


def test_fft_glue_fp16_operand_outputs():
    """Mode 2 (plain fp16 operand copy) of ftb_posenc / ftb_mha / ftb_mha_split equals fp16(mode 0) bit for bit."""
    gen = torch.Generator().manual_seed(11)
    B, L, E, heads = 2, 70, 256, 2
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    h = _handle()
    xin, _ = _fill(prog, 'xin', E, gen)
    pe = torch.randn(L + 5, E, generator=gen).to(DEV)
    scale = torch.tensor([0.7], device=DEV)
    p32, p16 = prog.buf('p32', E), prog.buf('p16', E, torch.float16)
    h.posenc(None, None, xin, pe, scale, E, B, L, g.stride, p32, round_tf32=False)
    h.posenc(None, None, xin, pe, scale, E, B, L, g.stride, p16, round_tf32=2)
    qkv, _ = _fill(prog, 'qkv', 3 * E, gen)
    a32, a16 = prog.buf('a32', E), prog.buf('a16', E, torch.float16)
    h.mha(qkv, E, heads, None, B, L, g.stride, a32, round_tf32=False)
    h.mha(qkv, E, heads, None, B, L, g.stride, a16, round_tf32=2)
    q3 = prog.buf('qkv16', 9 * E, torch.float16)
    q3.view(g.batch, g.stride, 9 * E)[:, :L] = _split3(prog.view('qkv').cpu()).to(DEV)
    s32, s16 = prog.buf('s32', E), prog.buf('s16', E, torch.float16)
    h.mha_split(q3, E, heads, None, B, L, g.stride, s32, round_tf32=False)
    h.mha_split(q3, E, heads, None, B, L, g.stride, s16, round_tf32=2)
    torch.cuda.synchronize()
    for a, b in (('p32', 'p16'), ('a32', 'a16'), ('s32', 's16')):
        assert torch.equal(prog.view(a).to(torch.float16), prog.view(b)), a
        assert float(prog.bufs[b].view(g.batch, g.stride, -1)[:, L:].abs().max()) == 0.0


@pytest.mark.parametrize('tag', ['zero_row', 'ties', 'all2'])
def test_lr_index_matches_reference_map(tag):
    """ftb_lr_index: the frame -> phoneme map of LengthRegulator.forward (common_layers.py:21-23) without the copy, on
    the reference goldens' durations (ties, zero durations, a zero-length sentence): bit-exact against the numpy oracle's
    idx (itself pinned to the reference's output on these files), and in its absolute-row form."""
    g = load_golden(f'lr_{tag}')
    h = _handle()
    dur = torch.from_numpy(g['dur']).to(DEV).contiguous()
    B, T = dur.shape
    _, idx_ref, lens_ref, _ = O.length_regulate_np(g['x'], g['dur'])
    want = torch.from_numpy(idx_ref).to(torch.int32)
    M = want.shape[1]
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(dur, B, T, False, cum, lens, meta)
    idx = torch.full((B, M), -7, dtype=torch.int32, device=DEV)
    h.lr_index(cum, lens, B, T, M, idx)
    torch.cuda.synchronize()
    assert torch.equal(idx.cpu(), want)
    h.lr_index(cum, lens, B, T, M, idx, row_stride=T + 8, pad_row=B * (T + 8))
    torch.cuda.synchronize()
    rows = torch.where(want >= 0, want + torch.arange(B, dtype=torch.int32)[:, None] * (T + 8),
                       torch.full_like(want, B * (T + 8)))
    assert torch.equal(idx.cpu(), rows)
    # gathering the source rows through the map reproduces the reference's LengthRegulator output bit for bit
    x = torch.from_numpy(g['x'])
    src = torch.cat([torch.nn.functional.pad(x, (0, 0, 0, 8)).reshape(B * (T + 8), -1), torch.zeros(1, x.shape[2])])
    assert np.array_equal(src[idx.cpu().long()].numpy().view(np.uint32), g['out'].view(np.uint32))


@pytest.mark.parametrize('cell,H,B,T', [('lstm', 512, 5, 23), ('lstm', 512, 40, 9), ('lstm', 512, 20, 17)])
def test_recurrence_reads_x_through_the_regulator_map(cell, H, B, T, monkeypatch):
    """ftb_rnn_problem.x_idx: the recurrence over the regulated sequence, its x-projection computed once per PHONEME and
    read through the frame -> row map (pad frames -> the b_ih row), is BIT-IDENTICAL to the recurrence over the
    materialised LengthRegulator output with a frame-level projection -- for both hand-off protocols."""
    gen = torch.Generator().manual_seed(3 * H + B)
    In = 64
    rnn = (torch.nn.GRU if cell == 'gru' else torch.nn.LSTM)(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    packed = pk.PackedRnn(rnn.to(DEV))
    h = _handle()
    dur = (torch.rand(B, T, generator=gen) * 4 - 0.5).to(DEV).contiguous()
    dur[B // 2, T // 2:] = 0.0                                         # a short sentence: padded frames behind it
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(dur, B, T, False, cum, lens, meta)
    M = int(meta.cpu()[0])
    xv = torch.randn(B, T, In, generator=gen).to(torch.float16)
    results = {}
    for mode in ('cluster', 'l2'):
        monkeypatch.setenv('FTB_RNN_HANDOFF', mode)
        # (a) per-phoneme projection + map
        tokp = eng.Program(h, eng.Geometry(B, T))
        a = tokp.buf('a', In, torch.float16)
        a.view(B, tokp.geo.stride, In)[:, :T] = xv.to(DEV)
        xproj = tokp.dense('xproj', (tokp.geo.cap_rows + 1, packed.ih.n), torch.float16)
        xproj[tokp.geo.cap_rows] = packed.ih.bias0.to(torch.float16)
        tokp.gemm(a, packed.ih, xproj, round_out=False)
        tokp.flush()
        tokp.run()
        fr = eng.Program(h, eng.Geometry(B, M))
        fr.dense('x_idx', (B, -1), torch.int32)
        h.lr_index(cum, lens, B, T, M, fr.bufs['x_idx'], row_stride=tokp.geo.stride, pad_row=tokp.geo.cap_rows)
        out_a = fr.buf('h', 2 * H, torch.float16)
        fr.rnn([(packed, xproj, out_a, 0, dict.__getitem__(fr.bufs, 'x_idx'))])
        fr.run()
        # (b) materialised regulator output + frame-level projection
        fr2 = eng.Program(h, eng.Geometry(B, M))
        lr = fr2.buf('lr', In, torch.float16)
        g = rt.LrGatherArgs()
        g.x, g.x_stride, g.x_ld, g.c1, g.c2 = a.data_ptr(), tokp.geo.stride, In, In, 0
        xf = a.float()                                                  # the gather reads fp32 rows
        g.x = xf.data_ptr()
        g.cum, g.lens, g.batch, g.T, g.M = cum.data_ptr(), lens.data_ptr(), B, T, M
        g.out, g.out_stride, g.out_ld, g.idx, g.round_tf32 = lr.data_ptr(), fr2.geo.stride, In, None, 2
        h.lr_gather(g)
        xp2 = fr2.buf('xp', packed.ih.n, torch.float16)
        out_b = fr2.buf('h', 2 * H, torch.float16)
        fr2.gemm(lr, packed.ih, xp2, round_out=False)
        fr2.rnn([(packed, xp2, out_b, 0)])
        fr2.run()
        torch.cuda.synchronize()
        assert fr.h.rnn_info(fr._rnn_plans[0])['handoff'] == mode
        assert torch.equal(fr.view('h'), fr2.view('h')), mode
        results[mode] = fr.view('h').clone()
        for pr in (tokp, fr, fr2):
            pr.close()
    assert torch.equal(results['cluster'], results['l2'])


def test_embed_and_heads_batched_match_single_and_torch():
    """ftb_embed_batch / ftb_rowdot_batch == the single-item entry points == torch (forward_tacotron.py:31,38-39)."""
    h = _handle()
    gen = torch.Generator().manual_seed(5)
    B, T = 3, 37
    geo = eng.Geometry(B, T)
    rows = B * geo.stride
    tok = torch.randint(0, 40, (B, T), generator=gen).to(DEV)
    tok2 = torch.randint(0, 3, (B, T), generator=gen).to(DEV)
    tabs = [torch.randn(40, e, generator=gen).to(DEV) for e in (64, 256, 32)]
    tab2 = torch.randn(3, 16, generator=gen).to(DEV)
    extra = torch.randn(B, 20, generator=gen).to(DEV)
    specs = [(tabs[0], None, None, None, True), (tabs[1], None, None, None, False),
             (tabs[2], tok2, tab2, extra, True), (tabs[1], None, None, None, 'f16')]
    singles, batched, items = [], [], []
    for table, t2, tb2, ex, mode in specs:
        width = table.shape[1] + (tb2.shape[1] if tb2 is not None else 0) + (ex.shape[1] if ex is not None else 0)
        dt = torch.float16 if mode == 'f16' else torch.float32
        a = torch.zeros(rows, width, device=DEV, dtype=dt)
        b = torch.zeros_like(a)
        h.embed(tok, table, a, width, B, T, geo.stride, tok2=t2, table2=tb2, extra=ex, round_tf32=bool(mode))
        items.append((tok, table, b, t2, tb2, ex, bool(mode)))
        singles.append(a)
        batched.append(b)
    h.embed_batch(items, B, T, geo.stride)
    torch.cuda.synchronize()
    for (table, t2, tb2, ex, mode), a, b in zip(specs, singles, batched):
        assert torch.equal(a, b)
        want = [table[tok]]
        if tb2 is not None:
            want.append(tb2[t2])
        if ex is not None:
            want.append(ex[:, None, :].expand(B, T, -1))
        want = torch.cat(want, dim=-1)
        got = b.view(B, geo.stride, -1)[:, :T]
        if mode == 'f16':
            assert torch.equal(got, want.half())
        elif mode:
            assert max_abs(got.cpu(), want.cpu()) <= 2 ** -11 * float(want.abs().max())
        else:
            assert torch.equal(got, want)
    # heads
    xs = [torch.randn(rows, k, generator=gen).to(DEV) for k in (128, 128, 64)]
    ws = [torch.randn(n, k, generator=gen).to(DEV) for n, k in ((1, 128), (1, 128), (4, 64))]
    bs = [torch.randn(w.shape[0], generator=gen).to(DEV) for w in ws]
    alphas = [1.3, 1.0, 1.0]
    ya = [torch.zeros(B, T, w.shape[0], device=DEV) for w in ws]
    yb = [torch.zeros_like(y) for y in ya]
    for x, w, b, al, y in zip(xs, ws, bs, alphas, ya):
        h.rowdot(x, x.shape[1], x.shape[1], w, b, w.shape[0], al, B, T, geo.stride, y)
    h.rowdot_batch([(x, w, b, w.shape[0], al, y) for x, w, b, al, y in zip(xs, ws, bs, alphas, yb)], B, T, geo.stride)
    torch.cuda.synchronize()
    for x, w, b, al, y1, y2 in zip(xs, ws, bs, alphas, ya, yb):
        assert torch.equal(y1, y2)
        want = (x.view(B, geo.stride, -1)[:, :T].double() @ w.double().t() + b.double()) / al
        assert max_abs(y2.double().cpu(), want.cpu()) < 1e-4
    with pytest.raises(RuntimeError):
        h.embed_batch(items * 3, B, T, geo.stride)                 # > FTB_MAX_GLUE_ITEMS


# --------------------------------------------------------------------------------------------------
# Split-precision path of the duration predictor (FTB_F_OUT_SPLIT, embed mode 3, ftb_rnn_problem.precise)
# --------------------------------------------------------------------------------------------------

def _split3(v):
    """[..., C] fp32 -> [..., 3*C8] fp16 [hi | lo | hi] (C8 = C rounded up to 8), the layout embed mode 3 / OUT_SPLIT write."""
    C = v.shape[-1]
    seg = pk.ceil_to(C, 8)
    hi = v.to(torch.float16)
    lo = (v - hi.float()).to(torch.float16)
    out = torch.zeros(v.shape[:-1] + (3 * seg,), dtype=torch.float16)
    out[..., :C], out[..., seg:seg + C], out[..., 2 * seg:2 * seg + C] = hi, lo, hi
    return out


@pytest.mark.parametrize('B,L,cin,cout,k', [(2, 50, 64, 256, 5), (3, 130, 256, 256, 5), (2, 40, 388, 256, 5),
                                            (2, 70, 256, 384, 1)])
def test_conv_split_precision(B, L, cin, cout, k):
    """[hi | lo | hi] operand x [W_hi | W_hi | W_lo] weight on kind::f16 MMAs: the conv at ~2^-20 relative operand
    precision on operands that are NOT pre-rounded (the single-fp16 product of the same operands is off by ~1e-3), result written back in the split layout."""
    gen = torch.Generator().manual_seed(B * 1000 + L + k + cin)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    v = torch.randn(B, L, cin, generator=gen)                       # NOT pre-rounded: the point of the split path
    seg = pk.ceil_to(cin, 8)
    a = prog.buf('a', 3 * seg, torch.float16)
    a.view(g.batch, g.stride, 3 * seg)[:, :L] = _split3(v).to(DEV)
    w = torch.randn(cout, cin, k, generator=gen) / (cin * k) ** 0.5
    gamma, beta = torch.rand(cout, generator=gen) + 0.5, torch.randn(cout, generator=gen) * 0.1
    mean, var = torch.randn(cout, generator=gen) * 0.1, torch.rand(cout, generator=gen) + 0.5
    packed = pk.pack_bn_conv(w.to(DEV), gamma.to(DEV), beta.to(DEV), mean.to(DEV), var.to(DEV), True, split=True)
    out = prog.buf('out', 3 * cout, torch.float16)
    out32 = prog.buf('out32', cout)
    prog.gemm(a, packed, out, split_in=True, split_out=True)
    prog.gemm(a, packed, out32, split_in=True, round_out=False)
    prog.flush()
    y = _conv_ref(v.double(), w, k).clamp_min(0)
    s = (gamma / torch.sqrt(var + 1e-5)).double()
    ref = y * s + (beta.double() - mean.double() * s)
    for impl in (1, 0):
        out.zero_(); out32.zero_()
        prog.h.impl = impl
        prog.run()
        torch.cuda.synchronize()
        got32 = prog.view('out32').cpu().double()
        o = prog.view('out').cpu()
        hi, lo, hi2 = o[..., :cout], o[..., cout:2 * cout], o[..., 2 * cout:]
        assert max_abs(got32, ref) < 1e-4, f'impl={impl}'        # fp32 accumulation over K = 3 * cin * k
        assert torch.equal(hi, hi2)
        assert max_abs(hi.double() + lo.double(), ref) < 1e-4, f'impl={impl}'
        assert torch.equal(hi, got32.float().to(torch.float16))     # hi is the rounded value, lo the remainder
    prog.h.impl = 0
    assert float(out.view(g.batch, g.stride, -1)[:, L:].float().abs().max()) == 0.0


def test_embed_split_mode():
    B, T, E, S = 3, 17, 12, 20
    gen = torch.Generator().manual_seed(5)
    table, table2 = torch.randn(30, E, generator=gen), torch.randn(4, 4, generator=gen)
    extra = torch.randn(B, S, generator=gen)
    tok, tok2 = torch.randint(0, 30, (B, T), generator=gen), torch.randint(0, 4, (B, T), generator=gen)
    prog = eng.Program(_handle(), eng.Geometry(B, T))
    seg = pk.ceil_to(E + 4 + S, 8)
    out = prog.buf('e', 3 * seg, torch.float16)
    prog.embed(tok.to(DEV), table.to(DEV), out, tok2=tok2.to(DEV), table2=table2.to(DEV), extra=extra.to(DEV),
               round_tf32=3)
    prog.run()
    torch.cuda.synchronize()
    v = torch.cat([table[tok], table2[tok2], extra[:, None, :].expand(-1, T, -1)], dim=2)
    assert torch.equal(prog.view('e').cpu(), _split3(v))


@pytest.mark.parametrize('H,B,L', [(64, 20, 60), (128, 9, 41), (64, 64, 200)])
def test_recurrence_precise(H, B, L):
    """ftb_rnn_problem.precise: split-fp16 recurrent product + ex2/rcp gates + unrounded fp32 h_t, fed by the split
    W_ih projection -- against the fp32 oracle GRU in float64-free plain fp32 (no operand rounding at all)."""
    gen = torch.Generator().manual_seed(H + B)
    In = 256
    rnn = torch.nn.GRU(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    packed = pk.PackedRnn(rnn.to(DEV), precise=True)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    v = torch.randn(B, L, In, generator=gen)
    a = prog.buf('x', 3 * In, torch.float16)
    a.view(g.batch, g.stride, 3 * In)[:, :L] = _split3(v).to(DEV)
    xp = prog.buf('xp', packed.ih.n)
    out = prog.buf('h', 2 * H)
    prog.gemm(a, packed.ih, xp, round_out=False, split_in=True)
    prog.rnn([(packed, xp, out, 0)])
    assert prog.info[-1]['handoff'] == 'cluster'
    prog.run()
    torch.cuda.synchronize()
    sd = {'r.' + k: v_.detach().cpu().double() for k, v_ in rnn.state_dict().items()}
    with torch.no_grad():
        ref = O.bi_rnn(sd, 'r', v.double(), 'gru')
    got = prog.view('h').cpu()
    assert max_abs(got, ref) < 2e-5, max_abs(got, ref)             # the single-fp16 recurrence sits at ~1e-3
    out.zero_()
    prog.run()
    torch.cuda.synchronize()
    assert torch.equal(prog.view('h').cpu(), got)


def test_program_set_seq_len_reuses_plans():
    """Launch-time geometry (ftb_*_plan_set_geometry): ONE program built for a capacity of 64 rows per sentence runs a
    k = 5 convolution + BiGRU at 50, 23, 64 and 50 rows again -- each time equal to a program built for exactly that
    length (bit-exact: same kernels, same tiles modulo the row count), with graphs replayed after the first run."""
    B, cin, cout, H = 5, 64, 96, 64
    gen = torch.Generator().manual_seed(77)
    w = torch.randn(cout, cin, 5, generator=gen) / (cin * 5) ** 0.5
    packed = pk.pack_conv(w.to(DEV), torch.randn(cout, generator=gen).to(DEV), relu=True)
    rnn = torch.nn.GRU(cout, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    prnn = pk.PackedRnn(rnn.to(DEV))

    def build(L):
        prog = eng.Program(_handle(), eng.Geometry(B, L))
        a = prog.buf('a', cin)
        y = prog.buf('y', cout)
        xp = prog.buf('xp', prnn.ih.n)
        h = prog.buf('h', 2 * H)
        mel = prog.dense('mel', (B, cout, -1))
        prog.gemm(a, packed, y, out_t=mel)
        prog.flush()
        prog.gemm(y, prnn.ih, xp, round_out=False)
        prog.rnn([(prnn, xp, h, 0)])
        return prog

    def fill(prog, v):
        g = prog.geo
        a = prog.bufs['a']
        a[:g.rows].view(g.batch, g.stride, cin)[:, :g.seq_len] = v.to(DEV)

    big = build(64)
    for L in (50, 23, 64, 50):
        v = pk.round_tf32(torch.randn(B, L, cin, generator=gen))
        big.set_seq_len(L)
        fill(big, v)
        for _ in range(2):                       # eager, then graph capture / replay
            big.run()
        exact = build(L)
        fill(exact, v)
        exact.run()
        torch.cuda.synchronize()
        for name in ('y', 'h'):
            assert torch.equal(big.view(name), exact.view(name)), (L, name)
        assert big.bufs['mel'].shape == (B, cout, L) and torch.equal(big.bufs['mel'], exact.bufs['mel'])
        g = big.geo
        for name in ('a', 'y', 'xp', 'h'):       # the gap rows of the current layout are zero
            b = big.bufs[name]
            assert float(b[:g.rows].view(B, g.stride, -1)[:, L:].float().abs().max()) == 0.0, (L, name)
        exact.close()
    with pytest.raises(AssertionError):
        big.set_seq_len(65)


@pytest.mark.parametrize('cell,H,B,mode', [('gru', 256, 33, 'cluster'), ('lstm', 512, 64, 'l2'), ('gru', 64, 17, 'cluster')])
def test_recurrence_repeatability_stress(cell, H, B, mode, monkeypatch):
    """Race evidence for the hand-off protocols (compute-sanitizer is closed on this pool, profiles/r2_sanitizer.txt): 20
    back-to-back runs of one plan -- eager launches first, CUDA-graph replays after -- are bit-identical.  A hand-off
    that is consumed early, lost or double-delivered changes h_t; the step count (150) gives it ~6000 chances per run."""
    monkeypatch.setenv('FTB_RNN_HANDOFF', mode)
    L, In = 150, 64
    gen = torch.Generator().manual_seed(H + B)
    rnn = (torch.nn.GRU if cell == 'gru' else torch.nn.LSTM)(In, H, batch_first=True, bidirectional=True)
    with torch.no_grad():
        for p in rnn.parameters():
            p.copy_((torch.rand(p.shape, generator=gen) * 2 - 1) / H ** 0.5)
    packed = pk.PackedRnn(rnn.to(DEV))
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    a, _ = _fill(prog, 'x', In, gen)
    xp = prog.buf('xp', packed.ih.n)
    out = prog.buf('h', 2 * H)
    prog.gemm(a, packed.ih, xp, round_out=False)
    prog.rnn([(packed, xp, out, 0)])
    assert prog.info[-1]['handoff'] == mode
    first = None
    for i in range(20):
        out.zero_()
        prog.run()
        torch.cuda.synchronize()
        got = prog.view('h').clone()
        assert bool(torch.isfinite(got).all())
        if first is None:
            first = got
        assert torch.equal(got, first), f'run {i} differs'


@pytest.mark.parametrize('B,L,cin,x16', [(3, 130, 256, False), (2, 300, 80, True), (1, 50, 256, True)])
def test_cbhg_tail_fused_matches_unfused_chain(B, L, cin, x16, monkeypatch):
    """K2f (ftb_cbhg_tail_*): pre_highway -> 4 highways -> W_ih projection in ONE kernel against (a) the same chain as six
    K1/K2 launches and (b) a float64 restatement of common_layers.py:118-123 on the same fp16 operands."""
    from forwardtacotron_b200.models.common_layers import CBHG
    torch.manual_seed(B * 100 + L)
    m = CBHG(K=2, in_channels=cin, channels=256, proj_channels=[256, cin], num_highways=4, dropout=0.0).to(DEV).eval()
    with torch.no_grad():
        for h in m.highways:
            h.W1.bias.normal_(0, 0.3)
            h.W2.bias.normal_(0, 0.3)
    from forwardtacotron_b200.models import _accel
    pack = _accel.pack_cbhg(m)
    gen = torch.Generator().manual_seed(L)
    v = torch.randn(B, L, cin, generator=gen).to(torch.float16)
    xdt = torch.float16 if x16 else torch.float32
    outs = {}
    for fused in (True, False):
        monkeypatch.setattr(eng, 'FUSED_TAIL', fused)
        prog = eng.Program(_handle(), eng.Geometry(B, L))
        g = prog.geo
        a = prog.buf('a', cin, torch.float16)
        a.view(g.batch, g.stride, cin)[:, :L] = v.to(DEV)
        xp = prog.buf('xp', pack.rnn.ih.n, xdt)
        if fused:
            prog.cbhg_tail(a, pack, xp)
        else:
            hw = [prog.buf('h0', 256, torch.float16), prog.buf('h1', 256, torch.float16)]
            hwf = [prog.buf('h0f', 256), prog.buf('h1f', 256)]
            prog.gemm(a, pack.pre_highway, hw[0], out_full=hwf[0])
            prog.flush()
            for i, w in enumerate(pack.highways):
                prog.gemm(hw[i & 1], w, hw[(i + 1) & 1], epilogue=rt.EPI_HIGHWAY, residual=hwf[i & 1],
                          out_full=hwf[(i + 1) & 1])
                prog.flush()
            prog.gemm(hw[0], pack.rnn.ih, xp, round_out=False)
        for _ in range(2):
            prog.run()
        torch.cuda.synchronize()
        outs[fused] = prog.view('xp').float().cpu()
        assert float(xp.view(g.batch, g.stride, -1)[:, L:].float().abs().max()) == 0.0     # gap rows untouched
    # float64 restatement on the same operands
    x = v.double() @ pack.pre_highway.w16[0].double().cpu().t()
    for h in m.highways:
        xq = x.to(torch.float16).double()
        x1 = xq @ h.W1.weight.detach().half().double().cpu().t() + h.W1.bias.detach().double().cpu()
        x2 = xq @ h.W2.weight.detach().half().double().cpu().t() + h.W2.bias.detach().double().cpu()
        gte = torch.sigmoid(x2)
        x = gte * x1.clamp_min(0) + (1 - gte) * x
    w_ih = torch.cat([m.rnn.weight_ih_l0, m.rnn.weight_ih_l0_reverse]).detach().half().double().cpu()
    b_ih = torch.cat([m.rnn.bias_ih_l0, m.rnn.bias_ih_l0_reverse]).detach().double().cpu()
    ref = x.to(torch.float16).double() @ w_ih.t() + b_ih
    tol = 2e-3 if x16 else 5e-4
    assert max_abs(outs[True], ref) < tol * max(1.0, float(ref.abs().max())), max_abs(outs[True], ref)
    assert max_abs(outs[False], ref) < tol * max(1.0, float(ref.abs().max()))
    assert max_abs(outs[True], outs[False]) < tol * max(1.0, float(ref.abs().max()))


@pytest.mark.parametrize('D,heads,B,L,masked', [(64, 2, 3, 70, True), (128, 2, 2, 130, False), (128, 2, 2, 333, True),
                                                (64, 2, 1, 64, False)])
def test_mha_tensor_core_split_precision(D, heads, B, L, masked):
    """ftb_mha_split (mma.sync on hi / lo fp16 operand pairs, fp32 online softmax) against float64 attention on operands
    that are NOT pre-rounded: 3e-6 -- a single-fp16 product of the same operands is off by ~1e-3."""
    E = D * heads
    gen = torch.Generator().manual_seed(D + L)
    prog = eng.Program(_handle(), eng.Geometry(B, L))
    g = prog.geo
    qkv = torch.randn(B, L, 3 * E, generator=gen) * 1.5
    buf = prog.buf('qkv16', 9 * E, torch.float16)
    buf.view(g.batch, g.stride, 9 * E)[:, :L] = _split3(qkv).to(DEV)
    out = prog.buf('att', E)
    out3 = prog.buf('att3', 3 * E, torch.float16)
    tok = None
    if masked:
        tok = torch.randint(1, 50, (B, L), generator=gen)
        for b in range(B):
            tok[b, L - 3 * b - 2:] = 0
        tok = tok.to(DEV)
    h = _handle()
    h.mha_split(buf, E, heads, tok, B, L, g.stride, out, round_tf32=False)
    h.mha_split(buf, E, heads, tok, B, L, g.stride, out3, round_tf32=3)
    torch.cuda.synchronize()
    q, k, v = [t.double().view(B, L, heads, D).transpose(1, 2) for t in qkv.chunk(3, dim=-1)]
    s = q @ k.transpose(-1, -2) / D ** 0.5
    if masked:
        s = s.masked_fill((tok.cpu() == 0)[:, None, None, :], float('-inf'))
    ref = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, L, E)
    got = prog.view('att').cpu().double()
    assert max_abs(got, ref) < 3e-6 * max(1.0, float(ref.abs().max())), max_abs(got, ref)
    o3 = prog.view('att3').cpu()
    assert torch.equal(o3[..., :E], o3[..., 2 * E:])
    assert max_abs(o3[..., :E].double() + o3[..., E:2 * E].double(), ref) < 4e-6 * max(1.0, float(ref.abs().max()))
    assert float(out.view(g.batch, g.stride, -1)[:, L:].abs().max()) == 0.0


def test_length_regulator_projection_branch_vs_oracle():
    """K6 fused into K5's source read (a.pitch != null): x + ps * Conv1d(1->C,3)(pitch) + es * Conv1d(1->C,3)(energy)
    (forward_tacotron.py:214-220) with the speaker tail appended (multi_forward_tacotron.py:280-282), then the gather --
    against the oracle's cond_projection + LengthRegulator on the same tensors, plus the idx / pad-row contract."""
    B, T, C1, C2 = 5, 37, 64, 16
    gen = torch.Generator().manual_seed(12)
    x = torch.randn(B, T, C1, generator=gen)
    extra = torch.randn(B, C2, generator=gen)
    pitch, energy = torch.randn(B, T, generator=gen), torch.randn(B, T, generator=gen)
    C = C1 + C2
    sd = {'p.weight': torch.randn(C, 1, 3, generator=gen) * 0.3, 'p.bias': torch.randn(C, generator=gen) * 0.1,
          'e.weight': torch.randn(C, 1, 3, generator=gen) * 0.3, 'e.bias': torch.randn(C, generator=gen) * 0.1}
    ps, es = 0.7, 1.3
    dur = torch.rand(B, T, generator=gen) * 4 - 0.5
    h = _handle()
    dd = dur.to(DEV).contiguous()
    cum = torch.empty((B, T), dtype=torch.int32, device=DEV)
    lens = torch.empty((B,), dtype=torch.int32, device=DEV)
    meta = torch.empty((4,), dtype=torch.int32, device=DEV)
    h.lr_scan(dd, B, T, False, cum, lens, meta)
    M = int(meta.cpu()[0])
    out = torch.full((B, M, C), 7.0, device=DEV)
    idx = torch.empty((B, M), dtype=torch.int32, device=DEV)
    keep = [t.to(DEV).contiguous() for t in (x, extra, pitch, energy, sd['p.weight'].reshape(C, 3), sd['p.bias'],
                                             sd['e.weight'].reshape(C, 3), sd['e.bias'])]
    a = rt.LrGatherArgs()
    a.x, a.x_stride, a.x_ld, a.c1 = keep[0].data_ptr(), T, C1, C1
    a.extra, a.c2 = keep[1].data_ptr(), C2
    a.pitch, a.energy = keep[2].data_ptr(), keep[3].data_ptr()
    a.wp, a.bp, a.ps = keep[4].data_ptr(), keep[5].data_ptr(), ps
    a.we, a.be, a.es = keep[6].data_ptr(), keep[7].data_ptr(), es
    a.cum, a.lens, a.batch, a.T, a.M = cum.data_ptr(), lens.data_ptr(), B, T, M
    a.out, a.out_stride, a.out_ld, a.idx, a.round_tf32 = out.data_ptr(), M, C, idx.data_ptr(), 0
    h.lr_gather(a)
    torch.cuda.synchronize()
    src = torch.cat([x, extra[:, None, :].expand(-1, T, -1)], dim=2)
    src = src + O.cond_projection(sd, 'p', pitch[:, None, :]) * ps + O.cond_projection(sd, 'e', energy[:, None, :]) * es
    ref, idx_ref, lens_ref, _ = O.length_regulate_np(src.numpy(), dur.numpy())
    assert ref.shape == (B, M, C)
    assert np.array_equal(idx.cpu().numpy(), idx_ref) and np.array_equal(lens.cpu().numpy(), lens_ref)
    got = out.cpu().numpy()
    assert np.abs(got - ref).max() < 2e-6                     # fp32 adds in the reference's order; conv summation order may differ
    for b in range(B):
        assert np.all(got[b, lens_ref[b]:] == 0.0)
