"""Prints parity errors of every golden case (diagnostic; the assertions live in tests/)."""
import sys, torch
sys.path.insert(0, '.')
from forwardtacotron_b200.configs import default_config
from forwardtacotron_b200.models import init_tts_model
from tests.helpers import load_golden, synthetic_sd, rel_l2, max_abs
DEV = torch.device('cuda', 0)
def model(tag, sd):
    m = init_tts_model(default_config(tag)); m.load_state_dict(sd); return m.to(DEV).eval()
for name in ['ft_spread_b3_t40', 'ft_fallback_b2_t24', 'ft_alpha_b2_t30', 'ft_b1_t74']:
    g = load_golden(name)
    sd = synthetic_sd('forward_tacotron', int(g['seed']), bool(g['spread']), float(g['dur_bias']))
    pm, em = float(g['pitch_mul']), float(g['energy_mul'])
    o = model('forward_tacotron', sd).generate(torch.from_numpy(g['x']).to(DEV), alpha=float(g['alpha']),
                                               pitch_function=lambda p: p * pm, energy_function=lambda e: e * em)
    line = f"{name}: dur {max_abs(o['dur'].cpu(), g['dur']):.2e} pitch {max_abs(o['pitch'].cpu(), g['pitch']):.2e}"
    if o['mel'].shape == g['mel'].shape:
        for k in ('mel', 'mel_post'):
            line += f" {k} abs {max_abs(o[k].cpu(), g[k]):.2e} rel {rel_l2(o[k].cpu(), g[k]):.2e}"
    print(line)
for name in ['mft_b1_t20', 'mft_b3_t32']:
    g = load_golden(name)
    sd = synthetic_sd('multi_forward_tacotron', int(g['seed']))
    o = model('multi_forward_tacotron', sd).generate(torch.from_numpy(g['x']).to(DEV),
                                                     torch.from_numpy(g['speaker_emb']).to(DEV))
    line = f"{name}: dur {max_abs(o['dur'].cpu(), g['dur']):.2e}"
    if o['mel'].shape == g['mel'].shape:
        for k in ('mel', 'mel_post'):
            line += f" {k} abs {max_abs(o[k].cpu(), g[k]):.2e} rel {rel_l2(o[k].cpu(), g[k]):.2e}"
    print(line)
for name in ['fp_b2_t30', 'fp_pad_b3_t36']:
    g = load_golden(name)
    sd = synthetic_sd('fast_pitch', int(g['seed']))
    m = model('fast_pitch', sd)
    x = torch.from_numpy(g['x']).to(DEV)
    o = m.generate(x, alpha=float(g['alpha']))
    tok = m.program_for('tok', x.shape[0], x.shape[1])
    line = (f"{name}: dur {max_abs(o['dur'].cpu(), g['dur']):.2e} pitch {max_abs(o['pitch'].cpu(), g['pitch']):.2e} "
            f"durT rel {rel_l2(tok.view('dur.out').cpu(), g['tap_dur_transformer']):.2e} "
            f"prenet rel {rel_l2(tok.view('prenet.out').cpu(), g['tap_prenet']):.2e}")
    if o['mel'].shape == g['mel'].shape:
        line += f" mel abs {max_abs(o['mel'].cpu(), g['mel']):.2e} rel {rel_l2(o['mel'].cpu(), g['mel']):.2e}"
    print(line)
