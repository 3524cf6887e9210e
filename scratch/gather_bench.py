"""Where the multi-GPU e2e step goes: generate() alone, + gather to rank 0, + gather to all, and the raw NCCL
send/recv of one mel block.   torchrun --nproc-per-node N scratch/gather_bench.py"""
import os, sys, time
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from forwardtacotron_b200 import parallel as par

rank, world, lr = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
dev = torch.device('cuda', lr)
torch.cuda.set_device(dev)
dist.init_process_group('nccl', device_id=dev)
model = bench.build_model('fallback').to(dev)
x = bench.make_tokens(64, 200, rank).to(dev)
for _ in range(3):
    out = model.generate(x)
mel, lens = out['mel_post'], par.mel_lengths(out['dur'])


def timed(fn, n=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / n * 1e3
    t = torch.tensor([a.elapsed_time(b) / n, wall], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def p2p():
    if rank == 0:
        bufs = [torch.empty_like(mel) for _ in range(world - 1)]
        ops = [dist.P2POp(dist.irecv, bufs[r - 1], r) for r in range(1, world)]
    else:
        ops = [dist.P2POp(dist.isend, mel, 0)]
    for w in dist.batch_isend_irecv(ops):
        w.wait()


res = {
    'generate': timed(lambda: model.generate(x)),
    'p2p_block_to_rank0': timed(p2p),
    'gather_only_dst0': timed(lambda: par.gather_variable_mels(mel, lens, dst=0)),
    'gather_only_all': timed(lambda: par.gather_variable_mels(mel, lens)),
    'generate+gather_dst0': timed(lambda: par.gather_variable_mels(*(lambda o: (o['mel_post'], par.mel_lengths(o['dur'])))(model.generate(x)), dst=0)),
    'generate+gather_all': timed(lambda: par.gather_variable_mels(*(lambda o: (o['mel_post'], par.mel_lengths(o['dur'])))(model.generate(x)))),
}
if rank == 0:
    for k, (gpu_ms, wall_ms) in res.items():
        print(f'{k:24s} device {gpu_ms:7.3f} ms   wall {wall_ms:7.3f} ms   (world {world}, block {mel.numel() * 4 / 1e6:.1f} MB)')
dist.destroy_process_group()
