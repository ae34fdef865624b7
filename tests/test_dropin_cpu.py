"""Drop-in checks against the LIVE reference tree (authoring container only: skipped where /root/reference is
absent, e.g. on the GPU box).  Checkpoints written by the reference's own PixelNeRFNet.save_weights load into the
mirror with strict=True and the other way round (reference src/model/models.py:268-316: plain state_dict pickles
at <checkpoints_path>/<name>/pixel_nerf_latest), and the reference's shipped conf files parse with the built-in
HOCON reader to what its scripts read from them."""
import os
import subprocess
import sys
import types

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="needs the reference tree")

_REF_SCRIPT = r'''
import sys, types, torch
sys.path.insert(0, {root!r})
sys.path.insert(0, {root!r} + "/tests/golden")
import make_golden as mg
ConfigTree, make_model, NeRFRenderer = mg.load_reference()
conf = mg.reference_conf(ConfigTree)
torch.manual_seed(7)
net = make_model(conf["model"])
args = types.SimpleNamespace(checkpoints_path={ckpt!r}, name="ref", resume=True)
mode = {mode!r}
if mode == "save":
    import os
    os.makedirs({ckpt!r} + "/ref", exist_ok=True)
    with torch.no_grad():
        for p in net.parameters():
            p.add_(torch.randn_like(p) * 0.01)       # not the init values: the load must really transfer them
    net.save_weights(args)
    print("CHECK", float(net.mlp_fine.blocks[3].fc_1.weight.double().sum()), float(net.encoder.model.layer3[2].conv2.weight.double().sum()))
else:
    args.name = "mirror"
    before = float(net.mlp_coarse.lin_z[1].weight.double().sum())
    net.load_weights(args, strict=True)
    print("CHECK", float(net.mlp_coarse.lin_z[1].weight.double().sum()), before)
'''


def _run_ref(mode, ckpt):
    res = subprocess.run([sys.executable, "-c", _REF_SCRIPT.format(root=ROOT, ckpt=str(ckpt), mode=mode)],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("CHECK")][-1]
    return [float(v) for v in line.split()[1:]]


def test_checkpoints_round_trip_with_the_live_reference(tmp_path):
    from pixelnerf_b200.model import make_model
    from pixelnerf_b200.testing import default_conf
    a, b = _run_ref("save", tmp_path)
    net = make_model(default_conf()["model"])
    args = types.SimpleNamespace(checkpoints_path=str(tmp_path), name="ref", resume=True)
    assert net.load_weights(args, strict=True, device="cpu") is net          # scripts chain on the return value
    assert abs(float(net.mlp_fine.blocks[3].fc_1.weight.double().sum()) - a) < 1e-9
    assert abs(float(net.encoder.model.layer3[2].conv2.weight.double().sum()) - b) < 1e-9
    # the other way: the mirror's save_weights, the reference's load_weights(strict=True)
    os.makedirs(tmp_path / "mirror")
    with torch.no_grad():
        net.mlp_coarse.lin_z[1].weight.add_(0.25)
    want = float(net.mlp_coarse.lin_z[1].weight.double().sum())
    args.name = "mirror"
    net.save_weights(args)
    assert os.path.exists(tmp_path / "mirror" / "pixel_nerf_latest")
    net.save_weights(args)                                                    # second save keeps a backup copy
    assert os.path.exists(tmp_path / "mirror" / "pixel_nerf_backup")
    got, before = _run_ref("load", tmp_path)
    assert abs(got - want) < 1e-6 and abs(got - before) > 1.0


@pytest.mark.parametrize("name", ["sn64", "srn", "dtu", "sn64_unseen", "pollen", "multi_obj"])
def test_shipped_confs_parse(name):
    from pixelnerf_b200.conf import ConfigFactory
    from pixelnerf_b200.render import NeRFRenderer
    conf = ConfigFactory.parse_file(os.path.join(REF, "conf", "exp", name + ".conf"))
    # what train/train.py and eval/*.py read (src/util/args.py, src/render/nerf.py:340-352, src/model/models.py:16-87)
    r = conf["renderer"]
    assert (r.get_int("n_coarse"), r.get_int("n_fine"), r.get_int("n_fine_depth")) == (64, 32, 16)
    assert abs(r.get_float("depth_std") - 0.01) < 1e-12 and r.get_list("sched", None) in ([], None)
    m = conf["model"]
    assert m.get_bool("use_encoder") and m.get_bool("use_xyz") and m.get_bool("use_viewdirs")
    assert m["mlp_coarse"].get_int("d_hidden") == 512 and m["mlp_coarse"].get_int("combine_layer") == 3
    assert m["code"].get_int("num_freqs") == 6 and abs(m["code"].get_float("freq_factor") - 1.5) < 1e-12
    assert m["encoder"].get_string("backbone") == "resnet34"
    if name.startswith("sn64"):
        assert m["encoder"].get_bool("use_first_pool") is False
    ren = NeRFRenderer.from_conf(r, lindisp=False, eval_batch_size=50000)
    assert (ren.n_coarse, ren.n_fine, ren.using_fine) == (64, 32, True)
