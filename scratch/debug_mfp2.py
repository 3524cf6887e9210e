import sys, torch
sys.path.insert(0, '.')
from tests.helpers import synthetic_sd
from forwardtacotron_b200 import synth
from forwardtacotron_b200.models import init_tts_model
from forwardtacotron_b200.configs import default_config
from oracle import ft_oracle as O
for seed in (11, 12, 13):
    sd = synthetic_sd('multi_fast_pitch', seed)
    m = init_tts_model(default_config('multi_fast_pitch')); m.load_state_dict(sd); m = m.cuda().eval()
    x = synth.synth_tokens(8, 48, seed=seed, with_pad=True)
    semb = synth.synth_speaker_emb(8, 256, seed=seed)
    ref = O.mfp_generate(sd, x, semb, alpha=0.9)
    out = m.generate(x.cuda(), semb.cuda(), alpha=0.9)
    torch.cuda.synchronize()
    got, r = out['dur'].cpu(), ref['dur']
    flips = (out['pitch_cond'].cpu() != ref['pitch_cond'])
    rd = lambda d: (d.clamp_min(0) + 0.5).long()
    diff = rd(got) != rd(r)
    print('seed', seed, 'flips', int(flips.sum()), 'maxabs', float((got - r).abs().max()), 'ndiff', int(diff.sum()))
    for b, t in diff.nonzero().tolist():
        print('   ', b, t, float(got[b, t]), float(r[b, t]), 'tok', int(x[b, t]))
    if not flips.any() and not diff.any():
        print('   mel rel', float((out['mel'].cpu() - ref['mel']).norm() / ref['mel'].norm()))
