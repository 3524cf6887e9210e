"""N-GPU check of parallel.generate_sharded over NCCL (torchrun): the gathered result equals the per-shard generate()
outputs, on every rank.   torchrun --nproc-per-node 2 scratch/check_sharded.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, '.')
from forwardtacotron_b200 import parallel as par, synth                 # noqa: E402
from forwardtacotron_b200.configs import default_config                # noqa: E402
from forwardtacotron_b200.models import init_tts_model                 # noqa: E402

rank, world, lr = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
dev = torch.device('cuda', lr)
torch.cuda.set_device(dev)
dist.init_process_group('nccl', device_id=dev)
model = init_tts_model(default_config('forward_tacotron'))
model.load_state_dict(synth.synth_state_dict(model.state_dict(), seed=0))
model = model.to(dev).eval()
x = synth.synth_tokens(2 * world + 1, 40, seed=3).to(dev)            # ragged split: the first rank gets one more
out = par.generate_sharded(model, x)
ok = True
for r in range(world):
    sl = par.shard_slice(x.shape[0], r, world)
    ref = model.generate(x[sl])
    lens = par.mel_lengths(ref['dur'])
    M = ref['mel_post'].shape[2]
    ok &= bool(torch.equal(out['mel_lens'][sl], lens))
    ok &= bool(torch.equal(out['mel_post'][sl, :, :M], ref['mel_post']))
    ok &= bool((out['mel_post'][sl, :, M:] == 0).all())
t = torch.tensor([int(ok)], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print('generate_sharded over NCCL, world', world, 'OK' if int(t) else 'MISMATCH', tuple(out['mel_post'].shape))
dist.destroy_process_group()
sys.exit(0 if int(t) else 1)
