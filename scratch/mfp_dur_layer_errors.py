"""Layer-by-layer error of the MultiFastPitch duration predictor (split-precision path) against a float64 oracle."""
import math, os, sys
import torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from forwardtacotron_b200 import synth
from forwardtacotron_b200.configs import default_config
from forwardtacotron_b200.models import init_tts_model
from oracle import ft_oracle as O
from tests.helpers import synthetic_sd

DEV = torch.device('cuda', 0)
seed = 11
sd = synthetic_sd('multi_fast_pitch', seed)
x = synth.synth_tokens(8, 48, seed=seed, with_pad=True); semb = synth.synth_speaker_emb(8, 256, seed=seed)
model = init_tts_model(default_config('multi_fast_pitch')); model.load_state_dict(sd); model = model.to(DEV).eval()
ref = O.mfp_generate(sd, x, semb, alpha=0.9)
out = model.generate(x.to(DEV), semb.to(DEV), alpha=0.9)
torch.cuda.synchronize()
print('dur err', float((out['dur'].cpu() - ref['dur']).abs().max()), 'mean', float((out['dur'].cpu() - ref['dur']).mean()))
st = model.__dict__['_accel']
B, T = x.shape
prog = model.program_for('tok', B, T)
prog.use_graph = False
sd64 = {k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()}
p = 'dur_pred.transformer'
e = F.embedding(x, sd64['dur_pred.embedding.weight'])
c = F.embedding(ref['pitch_cond'], sd64['dur_pred.conditional_embedding.weight'])
h = O._with_speaker(torch.cat([e, c], 2), semb.double())
h = h + sd64[p + '.pos_encoder.scale'] * sd64[p + '.pos_encoder.pe'][:T, 0, :][None]
E = h.shape[2]; heads = 2; d = E // heads
want = {}
want['x_full'] = h
for i in range(4):
    q = f'{p}.layers.{i}'
    qkv = F.linear(h, sd64[q + '.self_attn.in_proj_weight'], sd64[q + '.self_attn.in_proj_bias'])
    qq, kk, vv = [t.view(B, T, heads, d).transpose(1, 2) for t in qkv.chunk(3, dim=-1)]
    a = torch.softmax(torch.matmul(qq / math.sqrt(d), kk.transpose(-1, -2)), dim=-1)
    att = torch.matmul(a, vv).transpose(1, 2).reshape(B, T, E)
    y1 = h + F.linear(att, sd64[q + '.self_attn.out_proj.weight'], sd64[q + '.self_attn.out_proj.bias'])
    x1 = O.layer_norm(y1, sd64[q + '.norm1.weight'], sd64[q + '.norm1.bias'])
    c1 = F.relu(F.conv1d(x1.transpose(1, 2), sd64[q + '.conv1.weight'], sd64[q + '.conv1.bias'], padding=4))
    y2 = x1 + F.conv1d(c1, sd64[q + '.conv2.weight'], sd64[q + '.conv2.bias']).transpose(1, 2)
    h = O.layer_norm(y2, sd64[q + '.norm2.weight'], sd64[q + '.norm2.bias'])
    want[i] = dict(qkv=qkv, att=att, y1=y1, x1=x1, c1=c1.transpose(1, 2), y2=y2, x2=h)
final = O.layer_norm(h, sd64[p + '.norm.weight'], sd64[p + '.norm.bias'])

def split_val(buf, C):
    v = prog.view(buf).cpu()
    return v[..., :C].double() + v[..., C:2 * C].double()

def err(a, b):
    return f'max {float((a.double() - b).abs().max()):.2e} mean {float((a.double() - b).mean()):+.2e} (ref rms {float(b.pow(2).mean().sqrt()):.2f})'

prog.bufs['tok'].copy_(x.to(DEV)); prog.bufs['semb'].copy_(semb.to(DEV)); prog.alpha = 0.9
layer = {}
for info, step in zip(prog.info, prog.steps):
    step(); torch.cuda.synchronize()
    lab = info['label']
    if not lab.startswith('dur.'):
        continue
    if lab == 'dur.posenc':
        print(lab, err(prog.view('dur.x_full').cpu(), want['x_full']), '| split', err(split_val('dur.x', E), want['x_full']))
    for i in range(4):
        if lab == f'dur.l{i}.in_proj': print(lab, err(prog.view('dur.qkv').cpu(), want[i]['qkv']))
        if lab == f'dur.l{i}.out_proj': print(lab, err(prog.view('dur.blk_y').cpu(), want[i]['y1']))
        if lab == f'dur.l{i}.conv1': print(lab, err(split_val('dur.c1', want[i]['c1'].shape[2]), want[i]['c1']))
        if lab == f'dur.l{i}.conv2': print(lab, err(prog.view('dur.blk_y').cpu(), want[i]['y2']))
    if lab == 'dur.mha':
        n = layer.get('mha', 0); layer['mha'] = n + 1
        print(lab, n, err(split_val('dur.att', E), want[n]['att']))
    if lab == 'dur.norm1':
        n = layer.get('n1', 0); layer['n1'] = n + 1
        print(lab, n, err(prog.view('dur.x1_full').cpu(), want[n]['x1']), '| split', err(split_val('dur.x1', E), want[n]['x1']))
    if lab == 'dur.norm2':
        n = layer.get('n2', 0); layer['n2'] = n + 1
        if n < 3: print(lab, n, err(prog.view('dur.x2_full').cpu(), want[n]['x2']))
        else: print(lab, n, err(prog.view('dur.out').cpu(), final))
