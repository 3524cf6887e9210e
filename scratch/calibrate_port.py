"""Calibrates the timed CPU arm (the oracle PORT of generate(), bench.py --impl reference) against the REAL reference's
ForwardTacotron.generate() -- possible only in the build container, where /root/reference is importable (the GPU box has
no copy of it).  Same weights, same tokens, same thread count.   python scratch/calibrate_port.py [batch]"""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                                   # noqa: E402
from oracle import ft_oracle as O                              # noqa: E402

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 16
threads = os.cpu_count() or 1
torch.set_num_threads(threads)
ours = bench.build_model('fallback')
sd = {k: v.detach().clone() for k, v in ours.state_dict().items()}
x = bench.make_tokens(batch, 200)

sys.path.insert(0, '/root/reference')
for m in [k for k in sys.modules if k == 'models' or k.startswith('models.') or k == 'utils' or k.startswith('utils.')]:
    del sys.modules[m]
from models.forward_tacotron import ForwardTacotron            # noqa: E402  (the reference's own class)
from forwardtacotron_b200.configs import default_config        # noqa: E402
kw = dict(default_config('forward_tacotron')['forward_tacotron']['model'])
ref = ForwardTacotron(num_chars=135, n_mels=80, **kw).eval()
ref.load_state_dict(sd)


def once(fn):
    t0 = time.perf_counter()
    out = fn()
    return out, time.perf_counter() - t0


# The two programs ALTERNATE (port, reference, port, ...): timing one after the other in blocks picks up the drift of a
# shared host (a first version of this script reported 1.56x for what alternation shows to be 1.0x).
O.use_library_rnn(True)
tp, tr = [], []
with torch.no_grad():
    once(lambda: O.ft_generate(sd, x)); once(lambda: ref.generate(x))          # warm-up
    for _ in range(5):
        o_port, a = once(lambda: O.ft_generate(sd, x))
        o_ref, b = once(lambda: ref.generate(x))
        tp.append(a); tr.append(b)
O.use_library_rnn(False)
t_port, t_ref = sum(tp) / len(tp), sum(tr) / len(tr)
frames = int(o_port['mel_lens'].sum())
err = float((o_port['mel_post'] - o_ref['mel_post']).abs().max())
print(f'batch {batch} x 200, {threads} threads: reference generate() {t_ref * 1e3:.1f} ms ({frames / t_ref:.0f} frames/s), '
      f'oracle port {t_port * 1e3:.1f} ms ({frames / t_port:.0f} frames/s), port/reference time ratio {t_port / t_ref:.3f}, '
      f'mel_post max-abs difference {err:.2e}')
