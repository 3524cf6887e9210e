"""Sentence-sharded synthesis over NCCL on the GPUs of one box (BASELINE configs 4 and 5 are multi-GPU by definition):
parallel.generate_sharded == the oracle applied per shard.  Needs >= 2 GPUs (skipped otherwise); launched through
torch.distributed.run exactly like bench.py --gpus N.  The world-size-2 gloo tests of the same plumbing run on CPU
(tests/test_parallel.py)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gpus():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.timeout(900)
@pytest.mark.parametrize('case', ['small_ragged', 'config4_multift_128x200', 'config5_ft_32x4096'])
def test_generate_sharded_nccl_equals_per_shard_oracle(case):
    n = min(_gpus(), 8)
    if n < 2:
        pytest.skip(f'{n} GPU(s) visible: the sharded path needs at least 2 (gpurun --gpus N)')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={n}',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tests', 'sharded_worker.py'), case]
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=850)
    sys.stdout.write(res.stdout[-4000:])
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
