"""Shared test helpers: golden loading, synthetic state_dicts from the committed key tables."""
import json
import os

import numpy as np
import torch

from forwardtacotron_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
_DT = {'float32': torch.float32, 'int64': torch.int64}


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name + '.npz')) as z:
        return {k: z[k] for k in z.files}


def key_table(model_tag):
    with open(os.path.join(GOLDEN, f'state_dict_{model_tag}.json')) as f:
        return json.load(f)


def sinusoid_table(d_model, max_len=5000):
    """PositionalEncoding buffer (common_layers.py:132-140), shape [max_len,1,d_model]."""
    import math
    pe = torch.zeros(max_len, d_model)
    position = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
    pe[:, 0::2] = torch.sin(position * div)
    pe[:, 1::2] = torch.cos(position * div)
    return pe.unsqueeze(1)


def synthetic_sd(model_tag, seed, spread=True, dur_bias=None):
    """The same weights oracle/make_golden.py loaded into the reference, rebuilt from the
    committed key/shape table."""
    table = key_table(model_tag)
    template = {}
    for k, (shape, dt) in table.items():
        if k.endswith('pos_encoder.pe'):
            template[k] = sinusoid_table(shape[2], shape[0])
        else:
            template[k] = torch.empty(shape, dtype=_DT[dt])
    sd = synth.synth_state_dict(template, seed=seed, spread_durations=spread)
    if dur_bias is not None and not np.isnan(dur_bias):
        sd['dur_pred.lin.bias'] = torch.full_like(sd['dur_pred.lin.bias'], float(dur_bias))
    return sd


def rel_l2(a, b):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def max_abs(a, b):
    return float((torch.as_tensor(a, dtype=torch.float64) - torch.as_tensor(b, dtype=torch.float64)).abs().max())
