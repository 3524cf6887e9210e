import sys
import numpy as np, torch, torch.nn.functional as F
sys.path.insert(0, '.')
from tests.helpers import load_golden, synthetic_sd, rel_l2, max_abs
from forwardtacotron_b200.models import init_tts_model
from forwardtacotron_b200.configs import default_config
from oracle import ft_oracle as O
g = load_golden('mfp_b1_t24')
sd = synthetic_sd('multi_fast_pitch', int(g['seed']))
m = init_tts_model(default_config('multi_fast_pitch')); m.load_state_dict(sd); m = m.cuda().eval()
x = torch.from_numpy(g['x']); semb = torch.from_numpy(g['speaker_emb'])
out = m.generate(x.cuda(), semb.cuda(), alpha=float(g['alpha']))
torch.cuda.synchronize()
prog = m.__dict__['_accel']['programs'][('tok', 1, 24)]
cond = torch.from_numpy(g['pitch_cond'])
for p, tag, conditional in (('pitch_cond_pred', 'pitch_cond', False), ('energy_pred', 'energy', False),
                            ('dur_pred', 'dur', True), ('pitch_pred', 'pitch', True)):
    e = F.embedding(x, sd[p + '.embedding.weight'])
    if conditional:
        e = torch.cat([e, F.embedding(cond, sd[p + '.conditional_embedding.weight'])], dim=2)
    e = O._with_speaker(e, semb)
    print(tag, 'cat', max_abs(prog.view(f'{tag}.cat').cpu(), e))
    L = e.shape[1]
    pe = sd[p + '.transformer.pos_encoder.pe'][:L, 0, :]
    x0 = e + sd[p + '.transformer.pos_encoder.scale'] * pe[None]
    print(tag, 'x_full', max_abs(prog.view(f'{tag}.x_full').cpu(), x0))
    y = O.forward_transformer(sd, p + '.transformer', e, 2)
    print(tag, 'out rel', rel_l2(prog.view(f'{tag}.out').cpu(), y), 'E', e.shape[2])
    # first layer pieces
    lp = p + '.transformer.layers.0'
    att = O.self_attention(sd, lp + '.self_attn', x0, 2, None)
    x1 = O.layer_norm(x0 + att, sd[lp + '.norm1.weight'], sd[lp + '.norm1.bias'])
print('dur', out['dur'].cpu()[0, :6], g['dur'][0, :6])
