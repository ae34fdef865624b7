"""Shared test helpers: load golden fixtures and rebuild their inputs from seeds."""
import glob
import os

import numpy as np
import torch

from pixelnerf_b200 import synthetic as syn

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load(path):
    z = np.load(path, allow_pickle=False)
    return {k: z[k] for k in z.files}


def render_case_inputs(fx):
    """Rebuild weights / latent / draws of a render_*.npz fixture from its seed."""
    seed = int(fx["seed"])
    sb, ns = int(fx["sb"]), int(fx["ns"])
    shape = [int(v) for v in fx["latent_shape"]]
    n_rays = fx["rays"].shape[0] * fx["rays"].shape[1]
    return dict(
        wc=syn.mlp_weights(seed + 1),
        wf=syn.mlp_weights(seed + 2),
        latent=syn.latent_maps(seed + 5, sb * ns, shape[2], shape[3]),
        draws=syn.draws(seed + 6, n_rays, 64, 16, 16),
        cam=dict(poses=torch.from_numpy(fx["cam_poses"]), focal=torch.from_numpy(fx["cam_focal"]),
                 c=torch.from_numpy(fx["cam_c"]), image_shape=torch.from_numpy(fx["cam_image_shape"])),
        rays=torch.from_numpy(fx["rays"]),
        sb=sb, ns=ns, white=bool(fx["white"]), lindisp=bool(fx["lindisp"]),
    )
