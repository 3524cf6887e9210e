import sys, torch
sys.path.insert(0, '/root/repo')
from oracle import ft_oracle as O
from tests.helpers import load_golden, synthetic_sd, rel_l2, max_abs
orig_gru, orig_lstm = O.gru_direction, O.lstm_direction
orig_linear = O.linear
MODE = {'h': False}
def lin16(x, w, b=None):
    y = orig_linear(x, w, b)
    return y.half().float() if MODE['h'] else y
def wrap(fn):
    def f(x, w_ih, w_hh, b_ih, b_hh, reverse):
        O.linear = lin16
        try:
            return fn(x, w_ih, w_hh, b_ih, b_hh, reverse)
        finally:
            O.linear = orig_linear
    return f
O.gru_direction, O.lstm_direction = wrap(orig_gru), wrap(orig_lstm)
for name in ['ft_spread_b3_t40', 'ft_fallback_b2_t24', 'ft_alpha_b2_t30', 'ft_b1_t74']:
    g = load_golden(name)
    sd = synthetic_sd('forward_tacotron', int(g['seed']), bool(g['spread']), float(g['dur_bias']))
    pm, em = float(g['pitch_mul']), float(g['energy_mul'])
    x = torch.from_numpy(g['x'])
    outs = {}
    for h in (False, True):
        MODE['h'] = h
        outs[h] = O.ft_generate(sd, x, alpha=float(g['alpha']), pitch_function=lambda p: p * pm, energy_function=lambda e: e * em)
    a, b = outs[False], outs[True]
    if a['mel'].shape == b['mel'].shape:
        print(name, 'fp16-xproj vs base: mel rel', f"{rel_l2(b['mel'], a['mel']):.2e}", 'mel_post rel', f"{rel_l2(b['mel_post'], a['mel_post']):.2e}",
              '| base vs golden', f"{rel_l2(a['mel'], g['mel']):.2e}" if a['mel'].shape == g['mel'].shape else 'shape', '| fp16 vs golden',
              f"{rel_l2(b['mel'], g['mel']):.2e}" if b['mel'].shape == g['mel'].shape else 'shape', 'dur max', f"{max_abs(b['dur'], a['dur']):.1e}")
    else:
        print(name, 'shape changed', a['mel'].shape, b['mel'].shape)
