"""Pins oracle/ft_oracle.py to the reference: every oracle function is checked against
outputs the real reference produced (tests/golden/, made by oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import ft_oracle as O
from tests.helpers import load_golden, synthetic_sd, rel_l2, max_abs

TOL = 2e-5   # same ATen kernels; explicit-loop RNNs differ from the fused ones by ~1e-6


@pytest.mark.parametrize('tag', ['all2', 'ties', 'zero_row'])
def test_length_regulator_bit_exact(tag):
    g = load_golden(f'lr_{tag}')
    out, idx, lens, d = O.length_regulate_np(g['x'], g['dur'])
    assert out.shape == g['out'].shape
    assert np.array_equal(out.view(np.uint32), g['out'].view(np.uint32))
    assert np.array_equal(d.view(np.uint32), g['dur_after'].view(np.uint32))
    # idx is consistent with out
    for b in range(out.shape[0]):
        for m in range(out.shape[1]):
            want = g['x'][b, idx[b, m]] if idx[b, m] >= 0 else 0.0
            assert np.array_equal(out[b, m], want * np.ones_like(out[b, m]))


def test_cbhg_small():
    g = load_golden('cbhg_k5')
    import zlib  # noqa
    from forwardtacotron_b200 import synth
    # rebuild the module-local state_dict the generator used (keys have no model prefix)
    shapes = _cbhg_template(K=5, cin=24, ch=32, proj=(32, 24), nhw=2)
    sd = synth.synth_state_dict(shapes, seed=int(g['seed']), spread_durations=False)
    sd = {'c.' + k: v for k, v in sd.items()}
    with torch.no_grad():
        y = O.cbhg(sd, 'c', torch.from_numpy(g['x']))
    assert max_abs(y, g['y']) < TOL


def _cbhg_template(K, cin, ch, proj, nhw):
    t = {}
    def bnc(p, ci, co, k):
        t[p + '.conv.weight'] = torch.empty(co, ci, k)
        for n in ('weight', 'bias', 'running_mean', 'running_var'):
            t[f'{p}.bnorm.{n}'] = torch.empty(co)
        t[p + '.bnorm.num_batches_tracked'] = torch.empty((), dtype=torch.int64)
    for k in range(1, K + 1):
        bnc(f'conv1d_bank.{k - 1}', cin, ch, k)
    bnc('conv_project1', K * ch, proj[0], 3)
    bnc('conv_project2', proj[0], proj[1], 3)
    t['pre_highway.weight'] = torch.empty(ch, proj[1])
    for h in range(nhw):
        for w in ('W1', 'W2'):
            t[f'highways.{h}.{w}.weight'] = torch.empty(ch, ch)
            t[f'highways.{h}.{w}.bias'] = torch.empty(ch)
    for sfx in ('', '_reverse'):
        t['rnn.weight_ih_l0' + sfx] = torch.empty(3 * ch, ch)
        t['rnn.weight_hh_l0' + sfx] = torch.empty(3 * ch, ch)
        t['rnn.bias_ih_l0' + sfx] = torch.empty(3 * ch)
        t['rnn.bias_hh_l0' + sfx] = torch.empty(3 * ch)
    return t


@pytest.mark.parametrize('name', ['ft_spread_b3_t40', 'ft_fallback_b2_t24', 'ft_alpha_b2_t30', 'ft_b1_t74'])
@pytest.mark.parametrize('fused_rnn', [False, True])
def test_forward_tacotron_generate(name, fused_rnn):
    g = load_golden(name)
    sd = synthetic_sd('forward_tacotron', int(g['seed']), bool(g['spread']), float(g['dur_bias']))
    pm, em = float(g['pitch_mul']), float(g['energy_mul'])
    O.use_library_rnn(fused_rnn)
    try:
        taps = {}
        out = O.ft_generate(sd, torch.from_numpy(g['x']), alpha=float(g['alpha']),
                            pitch_function=lambda p: p * pm, energy_function=lambda e: e * em, taps=taps)
    finally:
        O.use_library_rnn(False)
    assert np.array_equal(out['dur'].numpy(), g['dur']) or max_abs(out['dur'], g['dur']) < TOL
    assert out['mel'].shape == g['mel'].shape
    for k in ('mel', 'mel_post', 'pitch', 'energy'):
        assert max_abs(out[k], g[k]) < 1e-4, k
        assert rel_l2(out[k], g[k]) < TOL, k
    if 'tap_prenet' in g:
        assert max_abs(taps['prenet'], g['tap_prenet']) < 1e-4
        assert np.array_equal(taps['lr'].numpy(), g['tap_lr']) or max_abs(taps['lr'], g['tap_lr']) < 1e-4
        assert max_abs(taps['lstm'], g['tap_lstm']) < 1e-4


@pytest.mark.parametrize('name', ['mft_b1_t20', 'mft_b3_t32'])
def test_multi_forward_tacotron_generate(name):
    g = load_golden(name)
    sd = synthetic_sd('multi_forward_tacotron', int(g['seed']))
    out = O.mft_generate(sd, torch.from_numpy(g['x']), torch.from_numpy(g['speaker_emb']))
    assert np.array_equal(out['pitch_cond'].numpy(), g['pitch_cond'])
    assert out['mel'].shape == g['mel'].shape
    for k in ('mel', 'mel_post', 'pitch', 'energy', 'dur'):
        assert max_abs(out[k], g[k]) < 1e-4, k


@pytest.mark.parametrize('name', ['fp_b2_t30', 'fp_pad_b3_t36'])
def test_fast_pitch_generate(name):
    g = load_golden(name)
    sd = synthetic_sd('fast_pitch', int(g['seed']))
    out = O.fp_generate(sd, torch.from_numpy(g['x']), alpha=float(g['alpha']))
    assert out['mel'].shape == g['mel'].shape
    for k in ('mel', 'pitch', 'energy', 'dur'):
        assert max_abs(out[k], g[k]) < 2e-4, k
        assert rel_l2(out[k], g[k]) < 5e-5, k


@pytest.mark.parametrize('name', ['mfp_b1_t24', 'mfp_pad_b3_t28'])
def test_multi_fast_pitch_generate(name):
    """MultiFastPitch oracle vs the reference (batch 1: unmodified reference; batch 3: the reference with the one
    batch-1-only arg-max line generalised, oracle/make_golden.py::BatchedMultiFP)."""
    g = load_golden(name)
    sd = synthetic_sd('multi_fast_pitch', int(g['seed']))
    out = O.mfp_generate(sd, torch.from_numpy(g['x']), torch.from_numpy(g['speaker_emb']), alpha=float(g['alpha']))
    assert out['mel'].shape == g['mel'].shape
    assert np.array_equal(out['pitch_cond'].numpy(), g['pitch_cond'])
    for k in ('mel', 'pitch', 'energy', 'dur'):
        assert max_abs(out[k], g[k]) < 2e-4, k
        assert rel_l2(out[k], g[k]) < 5e-5, k


def test_forward_tacotron_eval_forward():
    """forward(batch) in eval mode (teacher forcing + packed-sequence BiLSTM, forward_tacotron.py:118-165)."""
    from forwardtacotron_b200 import synth
    g = load_golden('ft_forward_b4_t26')
    sd = synthetic_sd('forward_tacotron', int(g['seed']))
    batch = synth.synth_batch(int(g['B']), int(g['T']), int(g['seed']))
    O.use_library_rnn(True)
    try:
        out = O.ft_forward(sd, batch)
    finally:
        O.use_library_rnn(False)
    for k in ('mel', 'mel_post', 'dur', 'pitch', 'energy'):
        assert out[k].shape == g[k].shape, k
        assert max_abs(out[k], g[k]) < 2e-4, k


def test_multi_forward_tacotron_eval_forward():
    """MultiForwardTacotron.forward(batch) in eval mode: predictors conditioned on the target pitch condition."""
    from forwardtacotron_b200 import synth
    g = load_golden('mft_forward_b4_t26')
    sd = synthetic_sd('multi_forward_tacotron', int(g['seed']))
    batch = synth.synth_batch(int(g['B']), int(g['T']), int(g['seed']), speaker_dims=256)
    O.use_library_rnn(True)
    try:
        out = O.mft_forward(sd, batch)
    finally:
        O.use_library_rnn(False)
    for k in ('mel', 'mel_post', 'dur', 'pitch', 'energy', 'pitch_cond'):
        assert out[k].shape == g[k].shape, k
        assert max_abs(out[k], g[k]) < 2e-4, k


def test_fast_pitch_eval_forward():
    """FastPitch.forward(batch) in eval mode (token mask on predictors + prenet, frame mask on the postnet)."""
    from forwardtacotron_b200 import synth
    g = load_golden('fp_forward_b4_t26')
    sd = synthetic_sd('fast_pitch', int(g['seed']))
    out = O.fp_forward(sd, synth.synth_batch(int(g['B']), int(g['T']), int(g['seed'])))
    for k in ('mel', 'dur', 'pitch', 'energy'):
        assert out[k].shape == g[k].shape, k
        assert max_abs(out[k], g[k]) < 2e-4, k


def test_multi_fast_pitch_eval_forward():
    """MultiFastPitch.forward(batch) in eval mode: token mask everywhere, target pitch condition, frame mask."""
    from forwardtacotron_b200 import synth
    g = load_golden('mfp_forward_b4_t26')
    sd = synthetic_sd('multi_fast_pitch', int(g['seed']))
    out = O.mfp_forward(sd, synth.synth_batch(int(g['B']), int(g['T']), int(g['seed']), speaker_dims=256))
    for k in ('mel', 'dur', 'pitch', 'energy', 'pitch_cond'):
        assert out[k].shape == g[k].shape, k
        assert max_abs(out[k], g[k]) < 2e-4, k
