"""Parity at the sizes BASELINE.json names (the benchmarked shapes), against the CPU oracle on the same weights and
tokens.  Every case has two parts:

  * mel / mel_post through the TEACHER-FORCED entry (`_generate_mel` fed the oracle's dur / pitch / energy), asserted at
    max-abs 1e-2 and relative L2 1e-3 unconditionally -- no dependence on duration ties;
  * a duration report from `generate()`: raw duration error, the number of rounded durations that differ from the
    oracle's, and how many of those lie outside the documented tie window -- asserted == 0.

The oracle runs in library-RNN mode (torch's fused CPU GRU/LSTM, the very kernels the reference calls); the largest case
takes a few seconds of CPU time.  scratch/scale_parity.py prints the same numbers (profiles/r2_err_report.txt)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from forwardtacotron_b200 import synth                                      # noqa: E402
from forwardtacotron_b200.configs import default_config                     # noqa: E402
from forwardtacotron_b200.models import init_tts_model                      # noqa: E402
from oracle import ft_oracle as O                                           # noqa: E402
from tests.helpers import synthetic_sd, rel_l2, max_abs                     # noqa: E402
from tests.test_gpu_models import DUR_WINDOW, DUR_WINDOW_MFP, MEL_ABS, MEL_REL, _rounded    # noqa: E402

DEV = torch.device('cuda', 0)
GEN = {'forward_tacotron': O.ft_generate, 'fast_pitch': O.fp_generate, 'multi_forward_tacotron': O.mft_generate,
       'multi_fast_pitch': O.mfp_generate}

CASES = [   # id, model, batch, T, alpha, spread durations, seed
    ('config2_ft_64x200_spread', 'forward_tacotron', 64, 200, 1.0, True, 21),
    ('config2_ft_64x200_fallback', 'forward_tacotron', 64, 200, 1.0, False, 22),      # the bench line: dur == 2, M = 400
    ('config3_fastpitch_64x200_spread', 'fast_pitch', 64, 200, 1.0, True, 23),
    ('config4_multift_128x200', 'multi_forward_tacotron', 128, 200, 1.0, True, 24),
    ('config4_multift_shard16x200', 'multi_forward_tacotron', 16, 200, 1.0, True, 25),   # one rank's share on 8 GPUs
    ('config5_ft_2x4096_alpha0.8', 'forward_tacotron', 2, 4096, 0.8, True, 26),
    ('multifp_16x200', 'multi_fast_pitch', 16, 200, 1.0, True, 11),
]


@pytest.mark.parametrize('name,tag,B,T,alpha,spread,seed', CASES, ids=[c[0] for c in CASES])
def test_parity_at_baseline_size(name, tag, B, T, alpha, spread, seed):
    sd = synthetic_sd(tag, seed, spread)
    x = synth.synth_tokens(B, T, seed=seed)
    multi = tag.startswith('multi')
    args = (x, synth.synth_speaker_emb(B, 256, seed=seed)) if multi else (x,)
    O.use_library_rnn(True)
    try:
        ref = GEN[tag](sd, *args, alpha=alpha)
    finally:
        O.use_library_rnn(False)
    model = init_tts_model(default_config(tag))
    model.load_state_dict(sd)
    model = model.to(DEV).eval()
    dargs = tuple(a.to(DEV) for a in args)

    # ---- part 1: mel via the teacher-forced entry, unconditional
    dur = ref['dur'].to(DEV).contiguous().clone()
    if multi:
        got = model._generate_mel(*dargs, dur, ref['pitch'].to(DEV), ref['pitch_cond'].to(DEV), ref['energy'].to(DEV))
    else:
        got = model._generate_mel(*dargs, dur, ref['pitch'].to(DEV), ref['energy'].to(DEV))
    torch.cuda.synchronize()
    for k in ('mel', 'mel_post'):
        assert got[k].shape == ref[k].shape, (k, got[k].shape, ref[k].shape)
        assert bool(torch.isfinite(got[k]).all()), k
        assert max_abs(got[k].cpu(), ref[k]) < MEL_ABS, k
        assert rel_l2(got[k].cpu(), ref[k]) < MEL_REL, k

    # ---- part 2: duration report of generate()
    out = model.generate(*dargs, alpha=alpha)
    torch.cuda.synchronize()
    d_got, d_ref = out['dur'].cpu(), ref['dur']
    if 'pitch_cond' in ref:
        assert torch.equal(out['pitch_cond'].cpu(), ref['pitch_cond']), 'pitch condition differs'
    window = DUR_WINDOW_MFP if tag == 'multi_fast_pitch' else DUR_WINDOW
    assert max_abs(d_got, d_ref) < window, max_abs(d_got, d_ref)
    flips = _rounded(d_got) != _rounded(d_ref)
    frac = (d_ref.clamp_min(0) + 0.5) % 1.0
    outside = flips & (torch.minimum(frac, 1 - frac) > window)
    assert int(outside.sum()) == 0, f'{int(outside.sum())} rounded durations differ outside the {window} tie window'
    assert max_abs(out['pitch'].cpu(), ref['pitch']) < 5e-3 and max_abs(out['energy'].cpu(), ref['energy']) < 5e-3
    if int(flips.sum()) == 0:          # then generate() itself must reproduce the oracle's mel at this size
        for k in ('mel', 'mel_post'):
            assert out[k].shape == ref[k].shape
            assert max_abs(out[k].cpu(), ref[k]) < MEL_ABS and rel_l2(out[k].cpu(), ref[k]) < MEL_REL, k
