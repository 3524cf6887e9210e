"""CPU-only checks: the C-ABI library loads and exports every declared symbol, the drop-in
state_dict contract holds, packing is exact, and the product path refuses to run without CUDA."""
import ctypes
import json
import os
import re

import pytest
import torch

from forwardtacotron_b200 import pack as pk
from forwardtacotron_b200 import runtime as rt
from forwardtacotron_b200.configs import default_config
from forwardtacotron_b200.models import init_tts_model
from oracle import ft_oracle as O
from tests.helpers import key_table, ROOT


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, 'include', 'ftb200.h')).read()
    declared = set(re.findall(r'\b(ftb_[a-z0-9_]+)\s*\(', header))
    assert declared == set(rt.SYMBOLS)
    lib = rt.load_library()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ftb_version() == 100
    assert b'aligned' in lib.ftb_error_string(-2)


def test_struct_layout_matches_header(tmp_path):
    """Compile the public header with gcc and compare sizeof/offsetof with the ctypes mirrors."""
    import subprocess
    src = tmp_path / 'probe.c'
    fields = {'ftb_gemm_problem': (rt.GemmProblem, ['w', 'bias0', 'out', 'out_t', 'seq_stride']),
              'ftb_rnn_problem': (rt.RnnProblem, ['xproj', 'out', 'out_col0', 'pad_value', 'out_f16', 'x_f16']),
              'ftb_lr_gather_args': (rt.LrGatherArgs, ['extra', 'wp', 'ps', 'we', 'cum', 'out', 'idx', 'round_tf32']),
              'ftb_embed_item': (rt.EmbedItem, ['table', 'e1', 'tok2', 'e2', 'extra', 'e3', 'out', 'out_ld', 'round_tf32']),
              'ftb_rowdot_item': (rt.RowdotItem, ['x_ld', 'k', 'w', 'bias', 'n_out', 'alpha', 'y'])}
    body = ''
    for cname, (_, fl) in fields.items():
        body += f'printf("%zu", sizeof({cname}));'
        for f in fl:
            body += f'printf(" %zu", offsetof({cname}, {f}));'
        body += 'printf("\\n");'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "ftb200.h"\nint main(){' + body + 'return 0;}')
    exe = tmp_path / 'probe'
    subprocess.check_call(['gcc', '-I', os.path.join(ROOT, 'include'), str(src), '-o', str(exe)])
    lines = subprocess.check_output([str(exe)], text=True).strip().splitlines()
    for line, (cname, (ct, fl)) in zip(lines, fields.items()):
        nums = [int(v) for v in line.split()]
        assert nums[0] == ctypes.sizeof(ct), cname
        for f, off in zip(fl, nums[1:]):
            assert getattr(ct, f).offset == off, (cname, f)


@pytest.mark.parametrize('tag', ['forward_tacotron', 'multi_forward_tacotron', 'fast_pitch', 'multi_fast_pitch'])
def test_state_dict_contract(tag):
    model = init_tts_model(default_config(tag))
    want = key_table(tag)
    got = {k: [list(v.shape), str(v.dtype).replace('torch.', '')] for k, v in model.state_dict().items()}
    assert list(got.keys()) == list(json.loads(json.dumps(want)).keys()) or set(got) == set(want)
    assert got == want


def test_round_tf32_matches_oracle_except_ties():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(100000, generator=g) * 3
    a, b = pk.round_tf32(x), O.round_tf32(x)
    assert float((a - x).abs().max() / x.abs().max()) < 2 ** -10
    assert (a.view(torch.int32) & 0x1FFF).eq(0).all()
    assert (a != b).sum() <= 30          # only exact ties differ (away-from-zero vs to-even)


def test_pack_highway_layout():
    u = 256
    w1, w2 = torch.randn(u, 64), torch.randn(u, 64)
    p = pk.pack_highway(w1, torch.zeros(u), w2, torch.ones(u), bn=256)
    assert p.w.shape == (1, 512, 64) and p.n == 512
    assert torch.equal(p.w[0, :128], pk.round_tf32(w1[:128]))
    assert torch.equal(p.w[0, 128:256], pk.round_tf32(w2[:128]))
    assert torch.equal(p.w[0, 256:384], pk.round_tf32(w1[128:]))
    assert torch.equal(p.bias0[:u], torch.zeros(u)) and torch.equal(p.bias0[u:], torch.ones(u))


def test_generate_refuses_cpu_tensors():
    model = init_tts_model(default_config('forward_tacotron'))
    with pytest.raises(RuntimeError, match='no CPU path'):
        model.generate(torch.ones(1, 10, dtype=torch.int64))


def test_training_forward_shapes():
    """Shape contract of the reference's own test (tests/test_forward_tacotron.py:19-38)."""
    model = init_tts_model(default_config('forward_tacotron')).train()
    batch = {'dur': torch.full((2, 10), 2.), 'mel': torch.ones(2, 80, 20), 'x': torch.ones(2, 10).long(),
             'mel_len': torch.full((2,), 20).long(), 'pitch': torch.ones(2, 10), 'energy': torch.ones(2, 10)}
    out = model(batch)
    assert out['mel'].shape == (2, 80, 20) and out['mel_post'].shape == (2, 80, 20)
    assert out['dur'].shape == (2, 10) and out['pitch'].shape == (2, 1, 10) and out['energy'].shape == (2, 1, 10)
    assert model.get_step() == 1


def test_torchscript_export_scripts_and_refuses_cpu(tmp_path):
    """forwardtacotron_b200.jit.script (README.md:159-171): the module scripts, its exported generate_jit is ONE call of
    the registered ftb200::generate_jit operator, the file round-trips, the rebuilt model aliases the flat weight buffer
    with an identical state_dict -- and a CPU call raises instead of falling back."""
    from forwardtacotron_b200 import jit as ftb_jit
    model = init_tts_model(default_config('forward_tacotron')).eval()
    scripted = ftb_jit.script(model)
    assert 'ftb200::generate_jit' in str(scripted.generate_jit.graph)
    path = str(tmp_path / 'tts.ts')
    torch.jit.save(scripted, path)
    loaded = torch.jit.load(path)
    rebuilt = ftb_jit._rebuild(loaded.flat, loaded.spec)
    sd0, sd1 = model.state_dict(), rebuilt.state_dict()
    assert list(sd0) == list(sd1) and all(torch.equal(sd0[k], sd1[k]) and sd0[k].dtype == sd1[k].dtype for k in sd0)
    lo, hi = loaded.flat.data_ptr(), loaded.flat.data_ptr() + loaded.flat.numel() * 4
    assert all(lo <= p.data_ptr() < hi for p in rebuilt.parameters())
    with pytest.raises(Exception, match='no CPU path'):
        loaded.generate_jit(torch.ones((1, 5), dtype=torch.int64))
    with pytest.raises(TypeError):
        ftb_jit.script(init_tts_model(default_config('fast_pitch')))


def test_misaligned_parameter_is_refused():
    """Packs alias some parameters and the kernels read them with 16-byte vector loads: a parameter that is a view at an
    odd offset of a larger buffer must be refused loudly (found by the TorchScript export, whose weights are views of one
    flat buffer -- it lays them out in 256-byte steps)."""
    model = init_tts_model(default_config('forward_tacotron')).eval()
    w = model.dur_pred.lin.weight
    big = torch.zeros(w.numel() + 1)
    big[1:] = w.detach().reshape(-1)
    model.dur_pred.lin.weight = torch.nn.Parameter(big[1:].view(w.shape), requires_grad=False)
    assert model.dur_pred.lin.weight.data_ptr() % 16 != 0
    with pytest.raises(RuntimeError, match='16-byte aligned'):
        model._accel_state(torch.device('cpu'))
