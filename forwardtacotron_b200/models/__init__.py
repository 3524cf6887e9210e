"""Drop-in `models` package surface (reference: models/__init__.py is empty; callers import the
submodules).  `init_tts_model` mirrors utils/checkpoints.py:37-49."""
from typing import Any, Dict


def init_tts_model(config: Dict[str, Any]):
    kind = config.get('tts_model', 'forward_tacotron')
    if kind == 'forward_tacotron':
        from .forward_tacotron import ForwardTacotron
        return ForwardTacotron.from_config(config)
    if kind == 'multi_forward_tacotron':
        from .multi_forward_tacotron import MultiForwardTacotron
        return MultiForwardTacotron.from_config(config)
    if kind == 'fast_pitch':
        from .fast_pitch import FastPitch
        return FastPitch.from_config(config)
    if kind == 'multi_fast_pitch':
        from .multi_fast_pitch import MultiFastPitch
        return MultiFastPitch.from_config(config)
    raise ValueError(f'Model type not supported: {kind}')
