"""Sanity + timing of the BASELINE.json configs that are not the bench headline (configs 3, 4, 5 and the 'spread'
regime of config 2) on one B200: runs generate(), checks finiteness / shapes / length bookkeeping, prints ms per call
and the per-launch device times.  Not a parity test (tests/test_gpu_models.py is); a scale test.

    python scratch/run_configs.py [fp] [mft] [long] [spread]
"""
import sys
import time

import torch

sys.path.insert(0, '.')
from forwardtacotron_b200.configs import default_config            # noqa: E402
from forwardtacotron_b200.models import init_tts_model              # noqa: E402
from forwardtacotron_b200 import parallel as par                    # noqa: E402

DEV = torch.device('cuda:0')


def tokens(batch, T, seed=1234, pad_tail=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(1, 135, (batch, T), generator=g, dtype=torch.int64)
    if pad_tail:                       # ragged batch: zero padding like the reference's collate
        for b in range(batch):
            n = T - (b * 7) % (T // 4)
            x[b, n:] = 0
    return x.to(DEV)


def timeit(fn, reps=5, warm=3):
    for _ in range(warm):
        out = fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3, out


def report(name, model, out, ms, T):
    mel = out['mel_post']
    lens = par.mel_lengths(out['dur'])
    ok = bool(torch.isfinite(mel).all()) and mel.shape[2] == int(lens.max())
    frames = int(lens.sum())
    print(f'{name}: mel_post {tuple(mel.shape)} finite+len ok={ok} frames {frames} {ms:.3f} ms/call '
          f'{frames / ms * 1e3 / 1e6:.2f} M frames/s', flush=True)
    st = model.__dict__.get('_accel')
    if st:
        rows = []
        for key, prog in st['programs'].items():
            for info, t in prog.run_timed():
                rows.append((t, key[0], info['label'], info['kind']))
        rows.sort(reverse=True)
        tot = sum(r[0] for r in rows)
        print(f'   launches {len(rows)} kernel sum {tot:.3f} ms; top: ' +
              '; '.join(f'{r[1]}:{r[2][:28]}={r[0]:.3f}' for r in rows[:8]), flush=True)
    assert ok, name


def main(which):
    torch.manual_seed(1234)
    if 'spread' in which:
        m = init_tts_model(default_config('forward_tacotron')).eval().to(DEV)
        with torch.no_grad():
            m.dur_pred.lin.weight.mul_(8.0)
            m.dur_pred.lin.bias.fill_(3.0)
        x = tokens(64, 200, pad_tail=True)
        ms, out = timeit(lambda: m.generate(x))
        report('config2-spread ForwardTacotron 64x200 ragged, spread durations', m, out, ms, 200)
    if 'fp' in which:
        m = init_tts_model(default_config('fast_pitch')).eval().to(DEV)
        x = tokens(64, 200)
        ms, out = timeit(lambda: m.generate(x))
        report('config3 FastPitch 64x200', m, out, ms, 200)
    if 'mft' in which:
        m = init_tts_model(default_config('multi_forward_tacotron')).eval().to(DEV)
        x = tokens(128, 200)
        g = torch.Generator().manual_seed(7)
        spk = torch.randn(128, 256, generator=g).to(DEV)
        ms, out = timeit(lambda: m.generate(x, spk))
        report('config4 MultiForwardTacotron 128x200 (1 GPU = the whole batch)', m, out, ms, 200)
    if 'long' in which:
        m = init_tts_model(default_config('forward_tacotron')).eval().to(DEV)
        x = tokens(32, 4096)
        ms, out = timeit(lambda: m.generate(x, alpha=0.8), reps=3, warm=2)
        report('config5 ForwardTacotron 32x4096 alpha=0.8', m, out, ms, 4096)


if __name__ == '__main__':
    main(sys.argv[1:] or ['spread', 'fp', 'mft', 'long'])
