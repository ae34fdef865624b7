"""Point features, fp32 ResnetFC and PixelNeRFNet.forward on the GPU vs the oracle and the
golden vectors of the live reference.  Tolerance: 1e-4 max-abs (north_star, fp32 path)."""
import numpy as np
import pytest
import torch

from oracle import model as omodel
from helpers import golden_files, load, render_case_inputs

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4


def _net_for(fx, inp, precision="fp32"):
    from pixelnerf_b200.testing import default_conf, load_mlp_weights, set_encoded
    from pixelnerf_b200.model import make_model
    from pixelnerf_b200.render import NeRFRenderer
    conf = default_conf(white_bkgd=inp["white"], mlp_precision=precision)
    net = make_model(conf["model"]).cuda().eval()
    load_mlp_weights(net, inp["wc"], inp["wf"])
    cam = inp["cam"]
    set_encoded(net, inp["latent"], cam["poses"], cam["focal"], cam["c"], cam["image_shape"], inp["ns"])
    ren = NeRFRenderer.from_conf(conf["renderer"], lindisp=inp["lindisp"]).cuda().eval()
    return net, ren


@pytest.mark.parametrize("path", golden_files("render_"))
def test_point_features_vs_oracle(path):
    from pixelnerf_b200 import _lib
    fx = load(path)
    inp = render_case_inputs(fx)
    net, _ = _net_for(fx, inp)
    xyz, dirs = torch.from_numpy(fx["probe_xyz"]), torch.from_numpy(fx["probe_dirs"])
    zx, _ = omodel.point_features(inp["latent"], inp["cam"], xyz, dirs, inp["ns"])
    fz, fxx = _lib.point_features(net._cam, net.latent_nhwc(torch.float32), xyz.shape[1], xyz=xyz.cuda(),
                                  viewdirs=dirs.cuda())
    assert (fz.cpu() - zx[:, :512]).abs().max() < 2e-5
    assert (fxx.cpu()[:, :42] - zx[:, 512:]).abs().max() < 2e-5
    assert float(fxx[:, 42:].abs().max()) == 0.0


@pytest.mark.parametrize("path", golden_files("render_"))
def test_net_forward_vs_golden(path):
    fx = load(path)
    inp = render_case_inputs(fx)
    net, _ = _net_for(fx, inp)
    xyz, dirs = torch.from_numpy(fx["probe_xyz"]).cuda(), torch.from_numpy(fx["probe_dirs"]).cuda()
    with torch.no_grad():
        out_c = net(xyz, coarse=True, viewdirs=dirs).cpu().numpy()
        out_f = net(xyz, coarse=False, viewdirs=dirs).cpu().numpy()
    # rgb within 1e-4 absolute; sigma (unbounded, O(10)) within 1e-4 relative to its scale
    for out, ref in ((out_c, fx["probe_coarse"]), (out_f, fx["probe_fine"])):
        assert np.abs(out[..., :3] - ref[..., :3]).max() < FP32_TOL
        assert np.abs(out[..., 3] - ref[..., 3]).max() < FP32_TOL * max(1.0, np.abs(ref[..., 3]).max())


def test_mlp_fp32_ragged_rows_and_views():
    """Row counts that are not multiples of the 128-row tile, NS = 1..3, vs the oracle MLP."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(42)
    packed = _lib.pack_weights_fp32({k: v.cuda() for k, v in w.items()})
    g = torch.Generator().manual_seed(0)
    for ns, ppo, sb in ((1, 1, 1), (1, 257, 1), (2, 130, 2), (3, 77, 1)):
        n = sb * ns * ppo
        zx = torch.cat([torch.rand(n, 512, generator=g), torch.randn(n, 42, generator=g)], -1)
        wd = dict(w)
        wd["_points_per_obj"] = ppo
        ref = omodel.resnetfc_forward(wd, zx, ns)
        fx = torch.nn.functional.pad(zx[:, 512:], (0, 2)).contiguous().cuda()
        out, pre = _lib.mlp_forward_fp32(packed, zx[:, :512].contiguous().cuda(), fx, sb, ns, ppo, want_pre=True)
        assert (pre.cpu() - ref).abs().max() < FP32_TOL * max(1.0, ref.abs().max().item())
        assert (out.cpu()[:, :3] - torch.sigmoid(ref[:, :3])).abs().max() < FP32_TOL


def test_unsupported_variants_raise():
    from pixelnerf_b200.testing import default_conf
    from pixelnerf_b200.model import make_model
    for path, val in (("model.mlp_coarse.type", "mlp"), ("model.mlp_coarse.combine_type", "max"),
                      ("model.mlp_coarse.d_hidden", 256), ("model.use_code_viewdirs", True),
                      ("model.use_global_encoder", True), ("model.encoder.backbone", "custom")):
        conf = default_conf()
        conf.put(path, val)
        with pytest.raises(NotImplementedError):
            make_model(conf["model"])
