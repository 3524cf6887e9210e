"""Latency of generate() at the small per-GPU batches the sharded BASELINE configs produce, on ONE GPU:
  1 x 74    one sentence (BASELINE config 1; the reference's README quotes "about 0.04 s" on a GeForce RTX 2080)
  4 x 4096  alpha 0.8: config 5's shard on 8 GPUs (a pure latency chain of 8192 recurrence steps)
  16 x 200  config 4's shard on 8 GPUs (MultiForwardTacotron is timed by bench.py --workload config4 itself)
usage: python scratch/small_batch_latency.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device('cuda', 0)
model = bench.build_model('fallback').to(dev)
for B, T, alpha in ((1, 74, 1.0), (4, 4096, 0.8), (16, 200, 1.0), (8, 200, 1.0)):
    x = bench.make_tokens(B, T).to(dev)
    for _ in range(4):
        out = model.generate(x, alpha=alpha)
    torch.cuda.synchronize()
    evs = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); out = model.generate(x, alpha=alpha); b.record(); evs.append((a, b))
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)[len(evs) // 2]
    frames = int((out['dur'] + 0.5).long().sum())
    prog = model.program_for('frame', B, out['mel'].shape[2])
    per = {i['label']: round(t, 4) for i, t in prog.run_timed() if i['kind'] == 'rnn'}
    print(f'{B} x {T} alpha {alpha}: {ms:.3f} ms per generate(), M = {out["mel"].shape[2]}, {frames / ms * 1e3:.3g} frames/s; '
          f'frame-side recurrences ms {per}', flush=True)
