"""pixelnerf_b200.install() against the LIVE reference tree (authoring container only; skipped where
/root/reference is absent).  install() must be an overlay: the reference's own ``util`` (with ``util.args``,
``bbox_sample``, ``cmap`` ...), ``model`` (with ``model.loss``) and ``render`` stay importable and only the
hot-path symbols are replaced.

* eval/gen_video.py is run UNMODIFIED (runpy) with a stub ``data`` package (one synthetic SRN-like object) and a
  stub ``imageio`` (absent from the image): it parses its arguments with ``util.args.parse_args``
  (src/util/args.py:9-112), builds the model / renderer, binds ``args.gpu_id``, generates the 40-pose ray tensor,
  encodes the source view and reaches its first ``render_par(rays[None])`` (eval/gen_video.py:63-217).  There is no
  GPU here and the product has no CPU fallback, so that call must raise the product's own RuntimeError -- everything
  before it must have worked.
* the train/train.py preamble (train/train.py:19-21, 72-93, 100-113 and the helpers its loop uses, :233-250) is
  replayed as a call sequence (the script itself imports matplotlib / a dataset at module level).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src")), reason="needs the reference tree")

_STUBS = r'''
import sys, types, torch, numpy as np
sys.path.insert(0, {root!r})

# --- stub dataset package: what eval/gen_video.py reads of it (get_split_dataset, dset[i], z_near, z_far, lindisp)
class _Dset:
    z_near, z_far, lindisp = 0.8, 1.8, False
    def __len__(self): return 1
    def __getitem__(self, i):
        g = torch.Generator().manual_seed(0)
        from pixelnerf_b200 import synthetic as syn
        poses = torch.stack([syn.pose_spherical(a, -10.0, 1.3) for a in (0.0, 30.0, 60.0)])
        return dict(path="synthetic/0", img_id=0, focal=torch.tensor(36.0), c=torch.tensor([16.0, 16.0]),
                    images=torch.rand(3, 3, 32, 32, generator=g) * 2 - 1, poses=poses)
data = types.ModuleType("data")
data.get_split_dataset = lambda fmt, datadir, want_split="all", training=True, **kw: (
    _Dset() if want_split != "all" else (_Dset(), _Dset(), _Dset()))
sys.modules["data"] = data
for name in ("imageio", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
'''

_GEN_VIDEO = _STUBS + r'''
import runpy, os
sys.path.insert(0, {ref!r} + "/src")          # what the script itself does first (eval/gen_video.py:4-6)
import pixelnerf_b200
pixelnerf_b200.install()                       # the two lines INTEGRATION.md asks a maintainer to add
import util, model, render
assert util.__file__.startswith({ref!r}), util.__file__          # still the reference's util ...
assert hasattr(util, "args") and hasattr(util, "bbox_sample") and hasattr(util, "cmap") and hasattr(util, "quat_to_rot")
assert hasattr(util, "get_image_to_tensor_balanced") and hasattr(util, "batched_index_select_nd")
from model import make_model, loss            # ... and the reference's model package with model.loss
assert make_model.__module__.startswith("pixelnerf_b200"), make_model.__module__
assert render.NeRFRenderer.__module__.startswith("pixelnerf_b200")
os.chdir({tmp!r})
# the reference's own conf, minus the ImageNet download (no network here): a HOCON include + one override
with open("srn_offline.conf", "w") as f:
    f.write('include required("' + {ref!r} + '/conf/exp/srn.conf")\nmodel {{ encoder {{ pretrained = False }} }}\n')
sys.argv = ["gen_video.py", "-n", "dropin", "-c", "srn_offline.conf", "-D", "unused", "--gpu_id", "0",
            "--num_views", "4", "-P", "1", "--split", "test", "-R", "512"]
try:
    runpy.run_path({ref!r} + "/eval/gen_video.py", run_name="__main__")
except RuntimeError as e:
    print("REACHED_RENDER", type(e).__name__, str(e)[:200])
else:
    raise SystemExit("gen_video.py finished without touching the renderer?")
'''

_TRAIN = _STUBS + r'''
import os
sys.path.insert(0, {ref!r} + "/src")
import pixelnerf_b200
pixelnerf_b200.install()
from model import make_model, loss             # train/train.py:19
from render import NeRFRenderer                # :20
from data import get_split_dataset             # :21
import util
from dotmap import DotMap                      # :33 (absent from the image: install() registered the stand-in)
os.chdir({tmp!r})
sys.argv = ["train.py", "-n", "dropin_train", "-c", {ref!r} + "/conf/exp/srn.conf", "-D", "unused", "--gpu_id", "0 1"]
def extra_args(parser):
    parser.add_argument("--batch_size", "-B", type=int, default=4)
    parser.add_argument("--nviews", "-V", type=str, default="2")
    parser.add_argument("--freeze_enc", action="store_true", default=None)
    return parser
args, conf = util.args.parse_args(extra_args, training=True, default_ray_batch_size=256)      # :72
assert args.gpu_id == [0, 1] and args.ray_batch_size == 256
device = util.get_cuda(args.gpu_id[0])
dset, val_dset, _ = get_split_dataset(args.dataset_format, args.datadir)
conf["model"]["encoder"].put("pretrained", False)     # no network for the ImageNet weights
net = make_model(conf["model"]).to(device=device)      # :80
net.stop_encoder_grad = args.freeze_enc
renderer = NeRFRenderer.from_conf(conf["renderer"], lindisp=dset.lindisp).to(device=device)   # :88
render_par = renderer.bind_parallel(net, args.gpu_id).eval()                                   # :93, two GPU ids
assert type(net).__module__.startswith("pixelnerf_b200") and net.mlp_precision == "fp16"
crit = loss.get_rgb_loss(conf["loss.rgb"], True)       # :110 (the reference's own model.loss)
assert conf.get_float("loss.lambda_coarse") == 1.0
# helpers the training loop calls on the reference's util (train/train.py:224-250)
d = dset[0]
H, W = d["images"].shape[-2:]
cam_rays = util.gen_rays(d["poses"], W, H, d["focal"], dset.z_near, dset.z_far, c=d["c"])      # CPU poses: reference code
assert cam_rays.shape == (3, H, W, 8)
pix = util.bbox_sample(torch.tensor([[4, 4, 20, 20.0]] * 3), 16)
assert pix.shape == (16, 3)
sel = util.batched_index_select_nd(d["images"].reshape(1, 3, -1).permute(0, 2, 1), torch.zeros(1, 5, dtype=torch.long))
assert sel.shape == (1, 5, 3)
assert abs(util.psnr(torch.zeros(4), torch.full((4,), 0.1)) - 20.0) < 1e-4
# the wrapper keeps the reference's contract (src/render/nerf.py:22-42): empty rays in, empty tensors out
out = render_par(torch.zeros(1, 0, 8).reshape(0, 0, 8))
assert out[0].shape == (0, 3)
sd = renderer.state_dict()                            # trainer saves it (train/train.py:118-135)
assert set(sd) == {{"iter_idx", "last_sched"}}
print("TRAIN_PREAMBLE_OK", type(render_par).__name__)
'''


def _run(script, tmp_path):
    res = subprocess.run([sys.executable, "-c", script.format(root=ROOT, ref=REF, tmp=str(tmp_path))],
                         capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, (res.stdout[-2000:], res.stderr[-4000:])
    return res.stdout


def test_gen_video_script_runs_unmodified_up_to_the_render(tmp_path):
    out = _run(_GEN_VIDEO, tmp_path)
    assert "Encoding source view(s)" in out and "Rendering" in out        # the script's own progress prints
    line = [l for l in out.splitlines() if l.startswith("REACHED_RENDER")][-1]
    assert "CUDA" in line and "no CPU fallback" in line, line            # the product refused, not something else
    assert os.path.isdir(tmp_path / "checkpoints" / "dropin")            # util.args.parse_args made its directories


def test_train_script_preamble_call_sequence(tmp_path):
    out = _run(_TRAIN, tmp_path)
    assert "TRAIN_PREAMBLE_OK" in out


def test_install_without_a_reference_tree_registers_the_mirrors(tmp_path):
    script = ("import sys; sys.path.insert(0, %r); import pixelnerf_b200; m, r, u = pixelnerf_b200.install(); "
              "import model, render, util, pyhocon; "
              "assert model is m and render.NeRFRenderer.__module__.startswith('pixelnerf_b200'); "
              "assert hasattr(util, 'gen_rays') and hasattr(pyhocon, 'ConfigFactory'); print('OK')" % ROOT)
    res = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, timeout=600, cwd=str(tmp_path))
    assert res.returncode == 0 and "OK" in res.stdout, res.stderr[-2000:]
