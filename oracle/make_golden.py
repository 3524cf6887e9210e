"""Generate tests/golden/*.npz from the REAL reference.  Run in the build container only
(needs /root/reference importable; the GPU box never runs this):

    python oracle/make_golden.py

Each fixture holds the inputs and the reference's outputs for key-derived synthetic
weights (forwardtacotron_b200/synth.py), so that the weights need not be stored: any box
rebuilds them from (key, shape, seed).  Also dumps the reference's state_dict key/shape
tables (the drop-in contract of SURVEY.md §8b) to tests/golden/state_dict_*.json.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get('FT_REFERENCE', '/root/reference')
sys.dont_write_bytecode = True
sys.path.insert(0, REF)
sys.path.insert(1, ROOT)

from models.common_layers import LengthRegulator, CBHG                     # noqa: E402  (reference)
from models.fast_pitch import FastPitch                                    # noqa: E402
from models.forward_tacotron import ForwardTacotron                        # noqa: E402
from models.multi_fast_pitch import MultiFastPitch                         # noqa: E402
from models.multi_forward_tacotron import MultiForwardTacotron             # noqa: E402
from utils.files import read_config                                        # noqa: E402

from forwardtacotron_b200 import synth                                     # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')
os.makedirs(OUT, exist_ok=True)
torch.set_num_threads(8)


def save(name, **arrays):
    arrays = {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in arrays.items()}
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **arrays)
    print(f'{name}: ' + ', '.join(f'{k}{list(v.shape)}' for k, v in arrays.items()))


def build(cls, yaml_name, seed, spread=True, dur_bias=None):
    cfg = read_config(os.path.join(REF, 'configs', yaml_name))
    model = cls.from_config(cfg).eval()
    sd = synth.synth_state_dict(model.state_dict(), seed=seed, spread_durations=spread)
    if dur_bias is not None:
        sd['dur_pred.lin.bias'] = torch.full_like(sd['dur_pred.lin.bias'], dur_bias)
    model.load_state_dict(sd)
    return model


def dump_keys():
    for cls, yaml_name, tag in [(ForwardTacotron, 'singlespeaker.yaml', 'forward_tacotron'),
                                (FastPitch, 'singlespeaker.yaml', 'fast_pitch'),
                                (MultiForwardTacotron, 'multispeaker.yaml', 'multi_forward_tacotron'),
                                (MultiFastPitch, 'multispeaker.yaml', 'multi_fast_pitch')]:
        cfg = read_config(os.path.join(REF, 'configs', yaml_name))
        sd = cls.from_config(cfg).state_dict()
        table = {k: [list(v.shape), str(v.dtype).replace('torch.', '')] for k, v in sd.items()}
        with open(os.path.join(OUT, f'state_dict_{tag}.json'), 'w') as f:
            json.dump(table, f, indent=0, sort_keys=True)
        print(tag, len(table), 'entries')


def with_taps(model, names):
    taps = {}
    hooks = []
    for n in names:
        mod = dict(model.named_modules())[n]
        def hook(m, i, o, n=n):
            taps[n] = (o[0] if isinstance(o, tuple) else o).detach().clone()
        hooks.append(mod.register_forward_hook(hook))
    return taps, hooks


def ft_case(name, B, T, seed, alpha=1.0, spread=True, dur_bias=None, pitch_mul=1.0, energy_mul=1.0, keep_taps=False):
    model = build(ForwardTacotron, 'singlespeaker.yaml', seed, spread, dur_bias)
    x = synth.synth_tokens(B, T, seed)
    taps, hooks = with_taps(model, ['prenet', 'lstm', 'lr', 'dur_pred.convs.0', 'dur_pred.rnn',
                                    'prenet.conv_project1', 'prenet.conv_project2', 'postnet'])
    out = model.generate(x, alpha=alpha, pitch_function=lambda p: p * pitch_mul,
                         energy_function=lambda e: e * energy_mul)
    for h in hooks:
        h.remove()
    save(name, x=x, alpha=np.float32(alpha), seed=seed, spread=int(spread),
         dur_bias=np.float32(np.nan if dur_bias is None else dur_bias),
         pitch_mul=np.float32(pitch_mul), energy_mul=np.float32(energy_mul),
         mel=out['mel'], mel_post=out['mel_post'], dur=out['dur'], pitch=out['pitch'], energy=out['energy'],
         **({f"tap_{k.replace('.', '_')}": v for k, v in taps.items()} if keep_taps else {}))


class BatchedMulti(MultiForwardTacotron):
    """The reference with the ONE B=1-only line (multi_forward_tacotron.py:252) replaced by a
    per-(b,t) arg-max; everything else is the reference's code."""

    def generate(self, x, speaker_emb, alpha=1.0, pitch_function=lambda x: x, energy_function=lambda x: x):
        self.eval()
        with torch.no_grad():
            pitch_cond_hat = self.pitch_cond_pred(x, speaker_emb)
            pitch_cond_hat = torch.argmax(pitch_cond_hat, dim=-1).long()
            dur_hat = self.dur_pred(x, pitch_cond_hat, speaker_emb, alpha=alpha).squeeze(-1)
            if torch.sum(dur_hat.long()) <= 0:
                torch.fill_(dur_hat, value=2.)
            pitch_hat = pitch_function(self.pitch_pred(x, pitch_cond_hat, speaker_emb).transpose(1, 2))
            energy_hat = energy_function(self.energy_pred(x, speaker_emb).transpose(1, 2))
            return self._generate_mel(x=x, dur_hat=dur_hat, pitch_hat=pitch_hat, energy_hat=energy_hat,
                                      pitch_cond_hat=pitch_cond_hat, semb=speaker_emb)


def mft_case(name, B, T, seed, patched):
    model = build(MultiForwardTacotron, 'multispeaker.yaml', seed)
    if patched:      # from_config returns the base class, so swap the class on the instance
        model.__class__ = BatchedMulti
    x = synth.synth_tokens(B, T, seed)
    semb = synth.synth_speaker_emb(B, 256, seed)
    out = model.generate(x, semb)
    save(name, x=x, speaker_emb=semb, seed=seed, patched=int(patched),
         mel=out['mel'], mel_post=out['mel_post'], dur=out['dur'], pitch=out['pitch'],
         energy=out['energy'], pitch_cond=out['pitch_cond'])


def ft_forward_case(name, B, T, seed):
    """ForwardTacotron.forward(batch) in eval mode (the GTA feature dump path, train_forward.py:33-51)."""
    model = build(ForwardTacotron, 'singlespeaker.yaml', seed)
    batch = synth.synth_batch(B, T, seed)
    out = model(batch)
    save(name, seed=seed, B=B, T=T, mel=out['mel'], mel_post=out['mel_post'], dur=out['dur'], pitch=out['pitch'],
         energy=out['energy'])


def mft_forward_case(name, B, T, seed):
    """MultiForwardTacotron.forward(batch) in eval mode (multi_forward_tacotron.py:186-241)."""
    model = build(MultiForwardTacotron, 'multispeaker.yaml', seed)
    batch = synth.synth_batch(B, T, seed, speaker_dims=256)
    out = model(batch)
    save(name, seed=seed, B=B, T=T, mel=out['mel'], mel_post=out['mel_post'], dur=out['dur'], pitch=out['pitch'],
         energy=out['energy'], pitch_cond=out['pitch_cond'])


def fp_forward_case(name, B, T, seed):
    """FastPitch.forward(batch) in eval mode (token / frame pad masks, fast_pitch.py:122-165)."""
    model = build(FastPitch, 'singlespeaker.yaml', seed)
    batch = synth.synth_batch(B, T, seed)
    out = model(batch)
    save(name, seed=seed, B=B, T=T, mel=out['mel'], dur=out['dur'], pitch=out['pitch'], energy=out['energy'])


def mfp_forward_case(name, B, T, seed):
    """MultiFastPitch.forward(batch) in eval mode (multi_fast_pitch.py:192-245)."""
    model = build(MultiFastPitch, 'multispeaker.yaml', seed)
    batch = synth.synth_batch(B, T, seed, speaker_dims=256)
    out = model(batch)
    save(name, seed=seed, B=B, T=T, mel=out['mel'], dur=out['dur'], pitch=out['pitch'], energy=out['energy'],
         pitch_cond=out['pitch_cond'])


class BatchedMultiFP(MultiFastPitch):
    """The reference with the ONE B=1-only line (multi_fast_pitch.py:243) replaced by a per-(b,t) arg-max;
    everything else is the reference's code."""

    def generate(self, x, speaker_emb, alpha=1.0, pitch_function=lambda x: x, energy_function=lambda x: x):
        self.eval()
        with torch.no_grad():
            pitch_cond_hat = self.pitch_cond_pred(x, speaker_emb=speaker_emb, alpha=alpha)
            pitch_cond_hat = torch.argmax(pitch_cond_hat, dim=-1).long()
            dur_hat = self.dur_pred(x, x_cond=pitch_cond_hat, speaker_emb=speaker_emb, alpha=alpha).squeeze(2)
            if torch.sum(dur_hat.long()) <= 0:
                torch.fill_(dur_hat, value=2.)
            pitch_hat = pitch_function(self.pitch_pred(x, x_cond=pitch_cond_hat, speaker_emb=speaker_emb).transpose(1, 2))
            energy_hat = energy_function(self.energy_pred(x, speaker_emb=speaker_emb).transpose(1, 2))
            return self._generate_mel(x=x, dur_hat=dur_hat, pitch_hat=pitch_hat, energy_hat=energy_hat,
                                      pitch_cond_hat=pitch_cond_hat, speaker_emb=speaker_emb)


def mfp_case(name, B, T, seed, patched, with_pad=False, alpha=1.0):
    model = build(MultiFastPitch, 'multispeaker.yaml', seed)
    if patched:
        model.__class__ = BatchedMultiFP
    x = synth.synth_tokens(B, T, seed, with_pad=with_pad)
    semb = synth.synth_speaker_emb(B, 256, seed)
    out = model.generate(x, semb, alpha=alpha)
    save(name, x=x, speaker_emb=semb, seed=seed, patched=int(patched), alpha=np.float32(alpha),
         with_pad=int(with_pad), mel=out['mel'], dur=out['dur'], pitch=out['pitch'], energy=out['energy'],
         pitch_cond=out['pitch_cond'])


def fp_case(name, B, T, seed, with_pad, alpha=1.0):
    model = build(FastPitch, 'singlespeaker.yaml', seed)
    x = synth.synth_tokens(B, T, seed, with_pad=with_pad)
    taps, hooks = with_taps(model, ['prenet', 'dur_pred.transformer', 'prenet.layers.0'])
    out = model.generate(x, alpha=alpha)
    for h in hooks:
        h.remove()
    save(name, x=x, alpha=np.float32(alpha), seed=seed, with_pad=int(with_pad),
         mel=out['mel'], dur=out['dur'], pitch=out['pitch'], energy=out['energy'],
         tap_prenet=taps['prenet'], tap_dur_transformer=taps['dur_pred.transformer'],
         tap_prenet_layer0=taps['prenet.layers.0'])


def lr_cases():
    g = torch.Generator().manual_seed(7)
    lr = LengthRegulator()
    # (i) all-2 fallback shape; (ii) U[-2,12) with exact .5 ties, zeros, negatives; (iii) an all-zero sentence
    x = torch.randn(4, 16, 8, generator=g)
    d1 = torch.full((4, 16), 2.0)
    d2 = torch.rand(4, 16, generator=g) * 14 - 2
    d2[0, :6] = torch.tensor([0.5, 1.5, 2.5, 0.49999997, 2.4999998, 0.0])
    d2[1, :4] = torch.tensor([-0.5, -3.0, 0.4999, 7.5])
    d3 = d2.clone()
    d3[2] = 0.0
    d3[3] = -1.0
    for tag, d in (('all2', d1), ('ties', d2), ('zero_row', d3)):
        din = d.clone()
        out = lr(x, din)          # clamps din in place (common_layers.py:18)
        save(f'lr_{tag}', x=x, dur=d, out=out, dur_after=din)


def cbhg_case():
    g = torch.Generator().manual_seed(11)
    m = CBHG(K=5, in_channels=24, channels=32, proj_channels=[32, 24], num_highways=2).eval()
    sd = synth.synth_state_dict(m.state_dict(), seed=3, spread_durations=False)
    m.load_state_dict(sd)
    x = torch.randn(2, 24, 19, generator=g)
    with torch.no_grad():
        y = m(x)
    save('cbhg_k5', x=x, y=y, seed=3)


if __name__ == '__main__':
    if sys.argv[1:] == ['ftfwd']:        # eval-mode forward(batch) fixtures only
        with torch.no_grad():
            ft_forward_case('ft_forward_b4_t26', 4, 26, seed=10)
        sys.exit(0)
    if sys.argv[1:] == ['mftfwd']:
        with torch.no_grad():
            mft_forward_case('mft_forward_b4_t26', 4, 26, seed=12)
        sys.exit(0)
    if sys.argv[1:] == ['mfpfwd']:
        with torch.no_grad():
            mfp_forward_case('mfp_forward_b4_t26', 4, 26, seed=13)
        sys.exit(0)
    if sys.argv[1:] == ['fpfwd']:
        with torch.no_grad():
            fp_forward_case('fp_forward_b4_t26', 4, 26, seed=11)
        sys.exit(0)
    if sys.argv[1:] == ['mfp']:          # add the MultiFastPitch fixtures without touching the existing ones
        with torch.no_grad():
            mfp_case('mfp_b1_t24', 1, 24, seed=8, patched=False)
            mfp_case('mfp_pad_b3_t28', 3, 28, seed=9, patched=True, with_pad=True, alpha=1.25)
        sys.exit(0)
    with torch.no_grad():
        dump_keys()
        lr_cases()
        cbhg_case()
        ft_case('ft_spread_b3_t40', 3, 40, seed=0, keep_taps=True)
        ft_case('ft_fallback_b2_t24', 2, 24, seed=1, spread=False, dur_bias=-1.0)
        ft_case('ft_alpha_b2_t30', 2, 30, seed=2, alpha=0.8, pitch_mul=1.5, energy_mul=0.5)
        ft_case('ft_b1_t74', 1, 74, seed=3)
        mft_case('mft_b1_t20', 1, 20, seed=4, patched=False)
        mft_case('mft_b3_t32', 3, 32, seed=5, patched=True)
        fp_case('fp_b2_t30', 2, 30, seed=6, with_pad=False)
        fp_case('fp_pad_b3_t36', 3, 36, seed=7, with_pad=True, alpha=1.25)
        mfp_case('mfp_b1_t24', 1, 24, seed=8, patched=False)
        mfp_case('mfp_pad_b3_t28', 3, 28, seed=9, patched=True, with_pad=True, alpha=1.25)
        ft_forward_case('ft_forward_b4_t26', 4, 26, seed=10)
        fp_forward_case('fp_forward_b4_t26', 4, 26, seed=11)
        mft_forward_case('mft_forward_b4_t26', 4, 26, seed=12)
        mfp_forward_case('mfp_forward_b4_t26', 4, 26, seed=13)
    with open(os.path.join(OUT, 'PROVENANCE.txt'), 'w') as f:
        f.write(f'generated by oracle/make_golden.py from {REF}\ntorch {torch.__version__}\n')
