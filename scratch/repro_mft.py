import sys, torch
sys.path.insert(0, '.')
from forwardtacotron_b200 import synth
from forwardtacotron_b200.configs import default_config
from forwardtacotron_b200.models import init_tts_model
from tests.helpers import synthetic_sd
DEV = torch.device('cuda', 0)
tag = sys.argv[1]
sd = synthetic_sd(tag, 5 if tag.startswith('multi') else 3)
m = init_tts_model(default_config(tag)); m.load_state_dict(sd); m = m.to(DEV).eval()
if tag.startswith('multi'):
    x = synth.synth_tokens(8, 64, seed=5); semb = synth.synth_speaker_emb(8, 256, seed=5)
    out = m.generate(x.to(DEV), semb.to(DEV), alpha=1.1)
else:
    x = synth.synth_tokens(2, 30, seed=3)
    out = m.generate(x.to(DEV))
torch.cuda.synchronize()
print({k: tuple(v.shape) for k, v in out.items()})
