"""Glue shared by the GPU tests, __graft_entry__.smoke() and bench.py: build a PixelNeRFNet on
a synthetic scene, and run the CPU oracle on the same scene.  The oracle import is local to
``oracle_render`` -- it is the checker / CPU baseline, never part of the product path."""
import torch

from .conf import ConfigTree
from .model import make_model
from .render import NeRFRenderer


def default_conf(white_bkgd=True, use_first_pool=True, mlp_precision="fp32"):
    """conf/default_mv.conf over conf/default.conf of the reference, encoder.pretrained off."""
    mlp = dict(type="resnet", n_blocks=5, d_hidden=512, combine_layer=3, combine_type="average")
    return ConfigTree.from_dict(dict(
        model=dict(use_encoder=True, use_global_encoder=False, use_xyz=True, canon_xyz=False, use_code=True,
                   code=dict(num_freqs=6, freq_factor=1.5, include_input=True), use_viewdirs=True,
                   use_code_viewdirs=False, mlp_coarse=dict(mlp), mlp_fine=dict(mlp), mlp_precision=mlp_precision,
                   encoder=dict(backbone="resnet34", pretrained=False, num_layers=4, use_first_pool=use_first_pool)),
        renderer=dict(n_coarse=64, n_fine=32, n_fine_depth=16, depth_std=0.01, sched=[], white_bkgd=white_bkgd)))


def set_encoded(net, latent, poses_w2c, focal, c, image_shape, ns):
    """Install an already-encoded scene (what encode() would cache) from explicit tensors."""
    dev = net.poses.device
    net.encoder.latent = latent.to(dev)
    ls = torch.tensor([latent.shape[-1], latent.shape[-2]], dtype=torch.float32)
    net.encoder.latent_scaling = (ls / (ls - 1) * 2.0).to(dev)
    net.poses, net.focal, net.c = poses_w2c.to(dev), focal.to(dev), c.to(dev)
    net.image_shape = image_shape.to(dev)
    net.num_views_per_obj = ns
    net.num_objs = latent.shape[0] // ns
    net._refresh_camera()


def load_mlp_weights(net, wc, wf):
    sd = net.state_dict()
    for prefix, w in (("mlp_coarse.", wc), ("mlp_fine.", wf)):
        for k, v in w.items():
            sd[prefix + k].copy_(v)


def scene_camera(sc):
    """Reference encode() camera handling (models.py:112-141) for a synthetic.scene dict."""
    poses = sc["src_poses"]
    rot = poses[:, :3, :3].transpose(1, 2)
    w2c = torch.cat((rot, -torch.bmm(rot, poses[:, :3, 3:])), dim=-1)
    f = sc["focal_t"]
    focal = (f[None, None].repeat(1, 2) if f.dim() == 0 else f.reshape(1, 2).clone()).float()
    focal[..., 1] *= -1.0
    image_shape = torch.tensor([float(sc["w"]), float(sc["h"])])
    c = (image_shape * 0.5).unsqueeze(0) if sc["c_t"] is None else sc["c_t"].reshape(1, 2)
    return dict(poses=w2c, focal=focal, c=c, image_shape=image_shape)


def build_net(sc, wc, wf, device, mlp_precision="fp32", lindisp=False):
    conf = default_conf(white_bkgd=sc["white"], mlp_precision=mlp_precision)
    net = make_model(conf["model"]).to(device).eval()
    load_mlp_weights(net, wc, wf)
    cam = scene_camera(sc)
    set_encoded(net, sc["latent"], cam["poses"], cam["focal"], cam["c"], cam["image_shape"], sc["ns"])
    renderer = NeRFRenderer.from_conf(conf["renderer"], lindisp=lindisp, eval_batch_size=50000).to(device).eval()
    return net, renderer


def oracle_render(sc, wc, wf, rays, draws, n_coarse=64, n_fine=32, n_fine_depth=16, lindisp=False, force=None):
    import oracle
    from oracle import model as omodel
    cam = scene_camera(sc)

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(wc if coarse else wf, sc["latent"], cam, sc["ns"], xyz, dirs)

    with torch.no_grad():
        return oracle.render_rays(net_fn, rays, 1, n_coarse, n_fine, n_fine_depth, draws, depth_std=0.01,
                                  white_bkgd=sc["white"], lindisp=lindisp, force=force)
