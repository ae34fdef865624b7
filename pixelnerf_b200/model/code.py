"""Positional encoding parameters (reference src/model/code.py).

The arithmetic runs inside pn_point_features (csrc/pn_features.cu); this module
exists so that configs, d_out bookkeeping and checkpoints (buffers ``_freqs`` and
``_phases``) stay identical to the reference.
"""
import numpy as np
import torch


class PositionalEncoding(torch.nn.Module):
    def __init__(self, num_freqs=6, d_in=3, freq_factor=np.pi, include_input=True):
        super().__init__()
        self.num_freqs, self.d_in, self.freq_factor = num_freqs, d_in, freq_factor
        self.include_input = include_input
        self.freqs = freq_factor * 2.0 ** torch.arange(0, num_freqs)
        self.d_out = num_freqs * 2 * d_in + (d_in if include_input else 0)
        self.register_buffer("_freqs", torch.repeat_interleave(self.freqs, 2).view(1, -1, 1))
        phases = torch.zeros(2 * num_freqs)
        phases[1::2] = np.pi * 0.5
        self.register_buffer("_phases", phases.view(1, -1, 1))

    def forward(self, x):
        raise NotImplementedError(
            "PositionalEncoding is fused into pn_point_features; call PixelNeRFNet.forward instead")

    @classmethod
    def from_conf(cls, conf, d_in=3):
        return cls(conf.get_int("num_freqs", 6), d_in, conf.get_float("freq_factor", np.pi),
                   conf.get_bool("include_input", True))
