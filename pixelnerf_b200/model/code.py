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
        """(N, 3) -> (N, 39) = [x, sin(f_k x), sin(f_k x + pi/2) ...].  Reference code.py:30-42.  The render path
        never calls this (the code is written by pn_point_features / pn_point_taps next to the latent lookup);
        direct callers get the same kernel through an identity camera and a dummy 2x2 latent."""
        from .. import _lib
        if self.num_freqs != 6 or abs(self.freq_factor - 1.5) > 1e-12 or not self.include_input or self.d_in != 3:
            raise NotImplementedError("the kernels implement num_freqs=6, freq_factor=1.5, include_input, d_in=3")
        if not x.is_cuda:
            raise RuntimeError("pixelnerf_b200 kernels need CUDA tensors (there is no CPU path)")
        dev = x.device
        n = x.shape[0]
        cam = dict(poses=torch.eye(3, 4, device=dev).unsqueeze(0).contiguous(), focal=torch.ones(1, 2, device=dev),
                   c=torch.zeros(1, 2, device=dev), scale_x=1.0, scale_y=1.0, n_objects=1, n_views=1)
        xyz = x.reshape(1, n, 3).float().contiguous()
        _, feat_x = _lib.point_features(cam, torch.zeros(1, 2, 2, _lib.D_LATENT, device=dev), n, xyz=xyz,
                                        viewdirs=torch.zeros_like(xyz), ldx=44)
        return feat_x[:, : self.d_out]

    @classmethod
    def from_conf(cls, conf, d_in=3):
        return cls(conf.get_int("num_freqs", 6), d_in, conf.get_float("freq_factor", np.pi),
                   conf.get_bool("include_input", True))
