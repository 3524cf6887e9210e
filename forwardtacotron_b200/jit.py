"""TorchScript export of `generate_jit` (reference: models/forward_tacotron.py:186-200, README.md:159-171
"Export Model with TorchScript": `model_script = torch.jit.script(tts_model); y = model_script.generate_jit(x)`).

The accelerated models reach their kernels through a C ABI, which `torch.jit.script` cannot compile.  What TorchScript
CAN call is an operator registered with the dispatcher, so this module registers

    ftb200::generate_jit(Tensor x, float alpha, float beta, Tensor flat, str spec) -> Tensor[]

and `script(model)` returns a scripted `nn.Module` whose exported `generate_jit(x, alpha, beta)` is one call of that
operator.  The module carries the weights itself: every floating-point entry of the `state_dict` lives in ONE flat fp32
buffer (it moves with `.to(device)` and is written by `torch.jit.save`), the constructor arguments, the key / shape /
offset table and the few integer buffers travel in the `spec` string.  The operator's CUDA implementation rebuilds the
accelerated model once per (spec, buffer) with its parameters ALIASING the buffer (`load_state_dict(assign=True)`) and
calls its `generate_jit` -- the same sm_100a kernels as `generate()`.  There is no CPU implementation (the operator
raises on CPU tensors, like the models do).

    from forwardtacotron_b200 import jit as ftb_jit
    model_script = ftb_jit.script(tts_model.cuda())            # instead of torch.jit.script(tts_model)
    y = model_script.generate_jit(x)                             # dict: mel, mel_post, dur, pitch, energy
    torch.jit.save(model_script, 'tts.ts')
    # later, in another process:  import forwardtacotron_b200.jit   (registers the operator), then torch.jit.load('tts.ts')
"""
import json
from typing import Dict, List

import torch
from torch import Tensor

_KEYS = ('mel', 'mel_post', 'dur', 'pitch', 'energy')
_LIB = torch.library.Library('ftb200', 'DEF')
_LIB.define('generate_jit(Tensor x, float alpha, float beta, Tensor flat, str spec) -> Tensor[]')

_MODELS: Dict[tuple, torch.nn.Module] = {}
_MAX_MODELS = 4
_ALIGN = 64                  # fp32 elements: 256 bytes


def _model_classes():
    from .models.forward_tacotron import ForwardTacotron
    return {'forward_tacotron': ForwardTacotron}


def _rebuild(flat: Tensor, spec: str) -> torch.nn.Module:
    info = json.loads(spec)
    model = _model_classes()[info['model']](**info['kwargs'])
    sd = {}
    for key, shape, off in info['float']:
        n = 1
        for s in shape:
            n *= s
        sd[key] = flat[off:off + n].view(shape)
    for key, shape, dtype, values in info['other']:
        sd[key] = torch.tensor(values, dtype=getattr(torch, dtype), device=flat.device).reshape(shape)
    model.load_state_dict(sd, strict=True, assign=True)        # parameters alias the flat buffer: no second copy in HBM
    for p in model.parameters():
        p.requires_grad_(False)
    return model.eval()


def _generate_jit_cuda(x: Tensor, alpha: float, beta: float, flat: Tensor, spec: str) -> List[Tensor]:
    if flat.device != x.device:
        raise RuntimeError(f'ftb200::generate_jit: x is on {x.device} but the weights are on {flat.device}')
    key = (spec, flat.data_ptr(), flat._version, x.device)
    model = _MODELS.get(key)
    if model is None:
        while len(_MODELS) >= _MAX_MODELS:
            _MODELS.pop(next(iter(_MODELS)))
        model = _MODELS[key] = _rebuild(flat, spec)
    out = model.generate_jit(x, alpha=alpha, beta=beta)
    return [out[k] for k in _KEYS]


def _generate_jit_cpu(x: Tensor, alpha: float, beta: float, flat: Tensor, spec: str) -> List[Tensor]:
    raise RuntimeError('ftb200::generate_jit: x and the module must be on a CUDA device -- this package has no CPU path '
                       '(the sm_100a kernels are the implementation)')


_LIB.impl('generate_jit', _generate_jit_cuda, 'CUDA')
_LIB.impl('generate_jit', _generate_jit_cpu, 'CPU')


class ScriptableTTS(torch.nn.Module):
    """What `script()` compiles: the flat weight buffer + the spec string; `generate_jit` is the operator call."""

    def __init__(self, flat: Tensor, spec: str):
        super().__init__()
        self.register_buffer('flat', flat)
        self.spec = spec

    @torch.jit.export
    def generate_jit(self, x: Tensor, alpha: float = 1.0, beta: float = 1.0) -> Dict[str, Tensor]:
        out = torch.ops.ftb200.generate_jit(x, alpha, beta, self.flat, self.spec)
        return {'mel': out[0], 'mel_post': out[1], 'dur': out[2], 'pitch': out[3], 'energy': out[4]}

    def forward(self, x: Tensor, alpha: float = 1.0, beta: float = 1.0) -> Dict[str, Tensor]:
        return self.generate_jit(x, alpha, beta)


def _ctor_kwargs(model) -> dict:
    """Constructor arguments of an accelerated model, recovered from its modules (what `from_config` splats from the yaml)."""
    import inspect
    kind = type(model).__name__
    if kind != 'ForwardTacotron':
        raise TypeError(f'script(): {kind} has no generate_jit in the reference (forward_tacotron.py:186 is the only one)')
    names = [p for p in inspect.signature(type(model).__init__).parameters if p != 'self']
    missing = [n for n in names if not hasattr(model, '_ctor_kwargs') or n not in model._ctor_kwargs]
    if missing:
        raise RuntimeError(f'script(): the model does not record its constructor arguments {missing}')
    return {n: model._ctor_kwargs[n] for n in names}


def script(model: torch.nn.Module) -> torch.jit.ScriptModule:
    """The scripted export of an accelerated ForwardTacotron (eval mode, like the reference's example)."""
    kwargs = _ctor_kwargs(model)
    floats, others, chunks, off = [], [], [], 0
    for key, t in model.state_dict().items():
        if t.is_floating_point():
            floats.append((key, list(t.shape), off))
            chunks.append(t.detach().to(torch.float32).reshape(-1))
            # every tensor starts on a 256-byte boundary of the buffer: the kernels read parameters with 16-byte vector
            # loads (embedding rows, predictor heads) and the packs may alias them
            pad = (-t.numel()) % _ALIGN
            if pad:
                chunks.append(torch.zeros(pad, dtype=torch.float32, device=t.device))
            off += t.numel() + pad
        else:
            others.append((key, list(t.shape), str(t.dtype).replace('torch.', ''), t.detach().cpu().reshape(-1).tolist()))
    flat = torch.cat(chunks).contiguous()
    spec = json.dumps({'model': 'forward_tacotron', 'kwargs': kwargs, 'float': floats, 'other': others}, sort_keys=True)
    return torch.jit.script(ScriptableTTS(flat, spec).eval())
