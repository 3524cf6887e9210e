"""Parity report at BASELINE sizes (profiles/r2_err_report.txt): for each config, generate() on the B200 vs the CPU oracle
on the same weights / tokens -- raw duration error, flipped rounded durations (and how many lie outside the documented
window), and mel / mel_post error through the teacher-forced entry fed with the ORACLE's dur / pitch / energy.
    python scratch/scale_parity.py [ft64 ft64f fp64 mft128 mft16 ft4096 mfp16 ...]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from forwardtacotron_b200 import synth                                      # noqa: E402
from forwardtacotron_b200.configs import default_config                     # noqa: E402
from forwardtacotron_b200.models import init_tts_model                      # noqa: E402
from oracle import ft_oracle as O                                           # noqa: E402
from tests.helpers import synthetic_sd, rel_l2, max_abs                     # noqa: E402

DEV = torch.device('cuda', 0)
CASES = {   # name: (model tag, batch, T, alpha, spread durations, seed)
    'ft64': ('forward_tacotron', 64, 200, 1.0, True, 21),
    'ft64f': ('forward_tacotron', 64, 200, 1.0, False, 22),
    'fp64': ('fast_pitch', 64, 200, 1.0, True, 23),
    'mft128': ('multi_forward_tacotron', 128, 200, 1.0, True, 24),
    'mft16': ('multi_forward_tacotron', 16, 200, 1.0, True, 25),
    'mfp16': ('multi_fast_pitch', 16, 200, 1.0, True, 11),
    'ft4096': ('forward_tacotron', 2, 4096, 0.8, True, 26),
}
GEN = {'forward_tacotron': O.ft_generate, 'fast_pitch': O.fp_generate, 'multi_forward_tacotron': O.mft_generate,
       'multi_fast_pitch': O.mfp_generate}


def rounded(d):
    return (torch.as_tensor(d).clamp_min(0) + 0.5).long()


def run(name):
    tag, B, T, alpha, spread, seed = CASES[name]
    sd = synthetic_sd(tag, seed, spread)
    x = synth.synth_tokens(B, T, seed=seed)
    multi = tag.startswith('multi')
    semb = synth.synth_speaker_emb(B, 256, seed=seed) if multi else None
    args = (x, semb) if multi else (x,)
    t0 = time.time()
    O.use_library_rnn(True)
    try:
        ref = GEN[tag](sd, *args, alpha=alpha)
    finally:
        O.use_library_rnn(False)
    t_cpu = time.time() - t0
    model = init_tts_model(default_config(tag))
    model.load_state_dict(sd)
    model = model.to(DEV).eval()
    dargs = tuple(a.to(DEV) for a in args)
    out = model.generate(*dargs, alpha=alpha)
    torch.cuda.synchronize()
    d_got, d_ref = out['dur'].cpu(), ref['dur']
    raw = max_abs(d_got, d_ref)
    flips = rounded(d_got) != rounded(d_ref)
    frac = (d_ref.clamp_min(0) + 0.5) % 1.0
    dist = torch.minimum(frac, 1 - frac)
    line = (f'{name:7s} {tag} {B}x{T} alpha={alpha} M_ref={ref["mel"].shape[2]} cpu={t_cpu:.1f}s | dur raw max-abs {raw:.2e} '
            f'flips {int(flips.sum())}/{flips.numel()} outside1e-3 {int((flips & (dist > 1e-3)).sum())} '
            f'outside2e-4 {int((flips & (dist > 2e-4)).sum())} max flip dist {float(dist[flips].max()) if flips.any() else 0:.1e}')
    if 'pitch_cond' in ref:
        line += f' | pitch_cond flips {int((out["pitch_cond"].cpu() != ref["pitch_cond"]).sum())}'
    line += f' | pitch max-abs {max_abs(out["pitch"].cpu(), ref["pitch"]):.1e} energy {max_abs(out["energy"].cpu(), ref["energy"]):.1e}'
    # mel through the teacher-forced entry with the oracle's own dur / pitch / energy
    dur_r = ref['dur'].to(DEV).contiguous().clone()
    if multi:     # (x, semb, dur, pitch, pitch_cond, energy)
        tf = dargs + (dur_r, ref['pitch'].to(DEV), ref['pitch_cond'].to(DEV), ref['energy'].to(DEV))
    else:
        tf = dargs + (dur_r, ref['pitch'].to(DEV), ref['energy'].to(DEV))
    got = model._generate_mel(*tf)
    torch.cuda.synchronize()
    for k in ('mel', 'mel_post'):
        assert got[k].shape == ref[k].shape, (got[k].shape, ref[k].shape)
        line += f' | {k} max-abs {max_abs(got[k].cpu(), ref[k]):.2e} rel-L2 {rel_l2(got[k].cpu(), ref[k]):.2e}'
    print(line, flush=True)


if __name__ == '__main__':
    for n in (sys.argv[1:] or list(CASES)):
        try:
            run(n)
        except Exception as e:                                               # keep the report going
            import traceback
            traceback.print_exc()
            print(f'{n}: FAILED {type(e).__name__}: {e}', flush=True)
