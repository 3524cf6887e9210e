import sys, torch
sys.path.insert(0, '.')
from tests.helpers import load_golden, synthetic_sd, rel_l2, max_abs
from forwardtacotron_b200 import synth
from forwardtacotron_b200.models import init_tts_model
from forwardtacotron_b200.configs import default_config
g = load_golden('ft_forward_b4_t26')
sd = synthetic_sd('forward_tacotron', int(g['seed']))
m = init_tts_model(default_config('forward_tacotron')); m.load_state_dict(sd); m = m.cuda().eval()
batch = {k: v.cuda() for k, v in synth.synth_batch(4, 26, 10).items()}
with torch.no_grad():
    out = m(batch)
ml = batch['mel_len'].cpu()
print('mel_len', ml.tolist(), 'target', g['mel'].shape)
for k in ('mel', 'mel_post'):
    got, want = out[k].cpu(), torch.from_numpy(g[k])
    for b in range(4):
        n = int(ml[b]); M = int(ml.max())
        print(k, b, 'valid abs', max_abs(got[b, :, :n], want[b, :, :n]), 'rel', rel_l2(got[b, :, :n], want[b, :, :n]),
              '| pad abs', max_abs(got[b, :, n:M], want[b, :, n:M]) if n < M else 0, 'pad mag', float(want[b, :, n:M].abs().max()) if n < M else 0)
