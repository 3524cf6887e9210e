"""Default model hyper-parameters for the mel-synthesis path.

These are plain data: the `<model>.model` sections of the reference's
configs/singlespeaker.yaml:98-128,152-187 and configs/multispeaker.yaml:100-138
(the kwargs that `Model.from_config` splats into the constructor,
forward_tacotron.py:242-247) so that tests, bench.py and smoke() can build the
named configurations on a box where the reference tree (and its yaml files) is
absent.  A user's own yaml still works through `Model.from_config(config)`.
"""
from copy import deepcopy

NUM_CHARS = 135      # len(utils.text.symbols.phonemes), utils/text/symbols.py:21-23
N_MELS = 80          # dsp.num_mels, configs/singlespeaker.yaml:12

_FORWARD_TACOTRON = dict(
    embed_dims=256, series_embed_dims=64,
    durpred_conv_dims=256, durpred_rnn_dims=64, durpred_dropout=0.5,
    pitch_conv_dims=256, pitch_rnn_dims=128, pitch_dropout=0.5, pitch_strength=1.0,
    energy_conv_dims=256, energy_rnn_dims=64, energy_dropout=0.5, energy_strength=1.0,
    prenet_dims=256, prenet_k=16, prenet_dropout=0.5, prenet_num_highways=4,
    rnn_dims=512,
    postnet_dims=256, postnet_k=8, postnet_num_highways=4, postnet_dropout=0.0,
)

_MULTI_FORWARD_TACOTRON = dict(
    speaker_emb_dims=256, embed_dims=256, series_embed_dims=128,
    durpred_conv_dims=256, durpred_rnn_dims=128, durpred_dropout=0.5,
    pitch_conv_dims=256, pitch_rnn_dims=256, pitch_dropout=0.5, pitch_strength=1.0,
    energy_conv_dims=256, energy_rnn_dims=64, energy_dropout=0.5, energy_strength=1.0,
    pitch_cond_conv_dims=256, pitch_cond_rnn_dims=128, pitch_cond_dropout=0.5,
    pitch_cond_emb_dims=4, pitch_cond_categorical_dims=3,
    prenet_dims=256, prenet_k=16, prenet_dropout=0.5, prenet_num_highways=4,
    rnn_dims=512,
    postnet_dims=256, postnet_k=8, postnet_num_highways=4, postnet_dropout=0.0,
)

_FAST_PITCH = dict(
    durpred_d_model=128, durpred_n_heads=2, durpred_layers=4, durpred_d_fft=128, durpred_dropout=0.5,
    pitch_d_model=128, pitch_n_heads=2, pitch_layers=4, pitch_d_fft=128, pitch_dropout=0.5,
    pitch_strength=1.0,
    energy_d_model=128, energy_n_heads=2, energy_layers=4, energy_d_fft=128, energy_dropout=0.5,
    energy_strength=1.0,
    d_model=256, conv1_kernel=9, conv2_kernel=1,
    prenet_layers=4, prenet_heads=2, prenet_fft=1024, prenet_dropout=0.1,
    postnet_layers=4, postnet_heads=2, postnet_fft=1024, postnet_dropout=0.1,
)


_MULTI_FAST_PITCH = dict(          # configs/multispeaker.yaml:165-216
    speaker_emb_dims=256,
    durpred_d_model=128, durpred_n_heads=2, durpred_layers=4, durpred_d_fft=128, durpred_dropout=0.5,
    pitch_d_model=128, pitch_n_heads=2, pitch_layers=4, pitch_d_fft=128, pitch_dropout=0.5,
    pitch_strength=1.0,
    energy_d_model=128, energy_n_heads=2, energy_layers=4, energy_d_fft=128, energy_dropout=0.5,
    energy_strength=1.0,
    pitch_cond_d_model=128, pitch_cond_n_heads=2, pitch_cond_layers=4, pitch_cond_d_fft=128,
    pitch_cond_dropout=0.5, pitch_cond_output_dims=3,
    d_model=256, conv1_kernel=9, conv2_kernel=1,
    prenet_layers=4, prenet_heads=2, prenet_fft=1024, prenet_dropout=0.1,
    postnet_layers=4, postnet_heads=2, postnet_fft=1024, postnet_dropout=0.1,
)


def default_config(tts_model: str = 'forward_tacotron') -> dict:
    """A config dict of the shape `read_config` (utils/files.py:22) returns,
    restricted to the keys the model constructors read."""
    table = {
        'forward_tacotron': _FORWARD_TACOTRON,
        'multi_forward_tacotron': _MULTI_FORWARD_TACOTRON,
        'fast_pitch': _FAST_PITCH,
        'multi_fast_pitch': _MULTI_FAST_PITCH,
    }
    if tts_model not in table:
        raise ValueError(f'Model type not supported: {tts_model}')
    return {
        'tts_model': tts_model,
        'dsp': {'num_mels': N_MELS},
        tts_model: {'model': deepcopy(table[tts_model])},
    }
