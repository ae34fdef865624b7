"""ResnetFC parameter container (reference src/model/resnetfc.py).

Holds the weights under the reference's state_dict names (``lin_in``, ``lin_out``,
``blocks.i.fc_{0,1}``, ``lin_z.i``) with the reference's initialisation; the
arithmetic is csrc/pn_mlp_fp32.cu (fp32) or csrc/pn_mlp_tc.cu (tensor cores).
Variants the kernels do not implement raise at construction instead of falling back.
"""
import torch
from torch import nn

from .. import _lib


class ResnetBlockFC(nn.Module):
    def __init__(self, size):
        super().__init__()
        self.size_in = self.size_h = self.size_out = size
        self.fc_0 = nn.Linear(size, size)
        self.fc_1 = nn.Linear(size, size)
        nn.init.constant_(self.fc_0.bias, 0.0)
        nn.init.kaiming_normal_(self.fc_0.weight, a=0, mode="fan_in")
        nn.init.constant_(self.fc_1.bias, 0.0)
        nn.init.zeros_(self.fc_1.weight)
        self.shortcut = None


class ResnetFC(nn.Module):
    def __init__(self, d_in, d_out=4, n_blocks=5, d_latent=0, d_hidden=128, beta=0.0, combine_layer=1000,
                 combine_type="average", use_spade=False):
        super().__init__()
        unsupported = []
        if d_hidden != _lib.D_HIDDEN: unsupported.append(f"d_hidden={d_hidden} (kernels: {_lib.D_HIDDEN})")
        if d_latent != _lib.D_LATENT: unsupported.append(f"d_latent={d_latent} (kernels: {_lib.D_LATENT})")
        if d_in != _lib.D_CODE: unsupported.append(f"d_in={d_in} (kernels: {_lib.D_CODE})")
        if d_out != 4: unsupported.append(f"d_out={d_out}")
        if n_blocks != _lib.N_BLOCKS: unsupported.append(f"n_blocks={n_blocks} (kernels: {_lib.N_BLOCKS})")
        if combine_layer != _lib.COMBINE_LAYER: unsupported.append(f"combine_layer={combine_layer}")
        if combine_type != "average": unsupported.append(f"combine_type={combine_type}")
        if beta > 0: unsupported.append("softplus activation (beta>0)")
        if use_spade: unsupported.append("use_spade")
        if unsupported:
            raise NotImplementedError("ResnetFC variant outside the B200 hot path: " + ", ".join(unsupported))
        self.lin_in = nn.Linear(d_in, d_hidden)
        self.lin_out = nn.Linear(d_hidden, d_out)
        for lin in (self.lin_in, self.lin_out):
            nn.init.constant_(lin.bias, 0.0)
            nn.init.kaiming_normal_(lin.weight, a=0, mode="fan_in")
        self.n_blocks, self.d_latent, self.d_in, self.d_out, self.d_hidden = n_blocks, d_latent, d_in, d_out, d_hidden
        self.combine_layer, self.combine_type, self.use_spade = combine_layer, combine_type, use_spade
        self.blocks = nn.ModuleList([ResnetBlockFC(d_hidden) for _ in range(n_blocks)])
        self.lin_z = nn.ModuleList([nn.Linear(d_latent, d_hidden) for _ in range(min(combine_layer, n_blocks))])
        for lin in self.lin_z:
            nn.init.constant_(lin.bias, 0.0)
            nn.init.kaiming_normal_(lin.weight, a=0, mode="fan_in")
        self.activation = nn.ReLU()
        self._packed = {}

    def invalidate(self):
        """Drop the kernel-side copies of the weights.  They are keyed on each parameter's (data_ptr, _version), which
        every autograd-visible write bumps (optimiser steps, ``load_state_dict``, ``copy_`` under ``no_grad``) --
        writes through ``p.data`` (some EMA / clipping code) do NOT: call this after such a write."""
        self._packed = {}

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self.invalidate()
        return out

    def _param_list(self):
        """[(state_dict name, module, attribute, parameter)] in ``state_dict()`` order, kept between calls and
        re-validated by identity against each module's ``_parameters`` (a re-assigned Parameter rebuilds it).
        ``state_dict()`` / ``parameters()`` cost 50-90 us per call -- four calls per rendered frame, fourteen per
        training step -- which was half of the host time of a frame."""
        plist = self.__dict__.get("_plist")
        if plist is not None:
            for _, m, n, p in plist:
                if m._parameters.get(n) is not p:
                    plist = None
                    break
        if plist is None:
            plist = [((prefix + "." if prefix else "") + n, m, n, p) for prefix, m in self.named_modules()
                     for n, p in m._parameters.items() if p is not None]
            self.__dict__["_plist"] = plist
        return plist

    def weight_dict(self):
        """{state_dict name: Parameter}, same keys and order as ``state_dict(keep_vars=True)``."""
        return {k: p for k, _, _, p in self._param_list()}

    def version_key(self):
        """(data_ptr, _version) of every parameter: the cache key of everything derived from the weights."""
        return tuple((p.data_ptr(), p._version) for _, _, _, p in self._param_list())

    def packed(self, precision):
        """Kernel-side view of the weights, rebuilt when any parameter changed."""
        key = self.version_key()
        hit = self._packed.get(precision)
        if hit is None or hit[0] != key:
            params = self.weight_dict()
            if precision == "fp32":
                hit = (key, _lib.pack_weights_fp32(params))
            elif precision == "fp16x3":       # precise / two-pass chains: fp16 hi + lo layers
                hit = (key, _lib.pack_weights_x3(params))
            elif precision == "bf16_train":   # row-major bf16 copies for the tensor-core backward
                hit = (key, _lib.pack_weights_bf16(params))
            else:
                hit = (key, _lib.pack_weights_tc(params, precision))
            self._packed[precision] = hit
        return hit[1]

    def forward(self, zx, combine_inner_dims=(1,), combine_index=None, dim_size=None):
        """(N, 512+42) rows ordered (object, view, point) -> raw lin_out outputs (N/NS, 4)."""
        assert zx.size(-1) == self.d_latent + self.d_in
        ns = combine_inner_dims[0] if len(combine_inner_dims) > 1 else 1
        ppo = combine_inner_dims[1] if len(combine_inner_dims) > 1 else zx.shape[0]
        zx = zx.reshape(-1, zx.shape[-1]).float()
        feat_z = zx[:, : self.d_latent].contiguous()
        feat_x = torch.nn.functional.pad(zx[:, self.d_latent:], (0, 2)).contiguous()
        n_obj = zx.shape[0] // (ns * ppo)
        _, pre = _lib.mlp_forward_fp32(self.packed("fp32"), feat_z, feat_x, n_obj, ns, ppo, want_pre=True)
        return pre

    @classmethod
    def from_conf(cls, conf, d_in, **kwargs):
        return cls(d_in, n_blocks=conf.get_int("n_blocks", 5), d_hidden=conf.get_int("d_hidden", 128),
                   beta=conf.get_float("beta", 0.0), combine_layer=conf.get_int("combine_layer", 1000),
                   combine_type=conf.get_string("combine_type", "average"),
                   use_spade=conf.get_bool("use_spade", False), **kwargs)
