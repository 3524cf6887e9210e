"""CPU study (oracle only): which operand roundings of the duration predictor move rounded durations at BASELINE
sizes.  Emulates the kernels' operand precision with oracle.set_operand_rounding.  python scratch/experiments/dur_precision_study.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import ft_oracle as O
from tests.helpers import synthetic_sd
from forwardtacotron_b200 import synth

f16 = lambda t: t.to(torch.float16).to(torch.float32)
torch.set_num_threads(8)
sd = synthetic_sd('forward_tacotron', seed=0)
B, T = int(os.environ.get('B', 64)), 200
x = synth.synth_tokens(B, T, seed=5)
emb = sd['dur_pred.embedding.weight'][x]


def dur(conv_q, rec_q):
    O.set_operand_rounding(conv_q, rec_q)
    try:
        return O.series_predictor(sd, 'dur_pred', emb)[:, :, 0]
    finally:
        O.set_operand_rounding(None, None)

ref = dur(None, None).double()
rr = torch.floor(ref.float() + 0.5)
print('dur range', float(ref.min()), float(ref.max()))
for name, cq, rq in [('fp16 conv+ih, fp16 rec', f16, f16), ('fp16 conv+ih, exact rec', f16, O.Identity),
                     ('exact conv+ih, fp16 rec', O.Identity, f16)]:
    d = dur(cq, rq)
    err = (d.double() - ref).abs()
    flips = torch.floor(d + 0.5) != rr
    frac = (ref + 0.5) - torch.floor(ref + 0.5)
    dist = torch.minimum(frac, 1 - frac)
    print(f'{name:28s} max-abs {float(err.max()):.2e} rms {float(err.pow(2).mean().sqrt()):.2e} flips {int(flips.sum())}'
          f' outside 1e-3: {int((flips & (dist > 1e-3)).sum())} outside 2e-4: {int((flips & (dist > 2e-4)).sum())}')
