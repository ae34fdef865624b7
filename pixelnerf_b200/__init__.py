"""pixelnerf_b200 -- B200 (sm_100a) implementation of pixelNeRF's render hot path.

Drop-in use from the reference's scripts (which do ``sys.path.insert(0, "src")`` and then
``from model import make_model; from render import NeRFRenderer; import util``):

    import pixelnerf_b200
    pixelnerf_b200.install()      # overlay: the reference's own util / model / render stay importable,
                                  # only the hot-path symbols are replaced

See INTEGRATION.md.
"""
import importlib
import os
import sys
import types

__version__ = "0.2.0"

_HERE = os.path.dirname(os.path.abspath(__file__))


def _third_party_shims():
    """``pyhocon`` and ``dotmap`` are imported by the reference (src/util/args.py:6, src/render/nerf.py:12) and
    are absent from this image: register the built-in stand-ins, but only when the real package does not import
    (a pyhocon that is installed and merely not imported yet is never shadowed)."""
    for name, make in (("pyhocon", _pyhocon_shim), ("dotmap", _dotmap_shim)):
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except ImportError:
            sys.modules[name] = make()


def _pyhocon_shim():
    from . import conf
    return conf


def _dotmap_shim():
    from .render.nerf import DotMap
    m = types.ModuleType("dotmap")
    m.DotMap = DotMap
    m.__doc__ = "pixelnerf_b200 stand-in for dotmap (attribute dict with auto-vivification and toDict)"
    return m


def _is_ours(mod):
    return os.path.abspath(getattr(mod, "__file__", "") or "").startswith(_HERE)


def _import_reference(name):
    """The reference's own top-level package ``name`` (found through sys.path, i.e. its ``src`` directory), or
    None when it is not importable / what is found is this package's mirror."""
    if name in sys.modules and not _is_ours(sys.modules[name]):
        return sys.modules[name]
    try:
        mod = importlib.import_module(name)
    except ImportError:
        return None
    return None if _is_ours(mod) else mod


def install(reference_src=None):
    """Route the reference's hot path onto this package.

    With the reference's ``src`` directory on ``sys.path`` (its scripts put it there; or pass it as
    ``reference_src``) this is an OVERLAY: ``util``, ``model`` and ``render`` remain the reference's own packages --
    ``util.args.parse_args``, ``util.bbox_sample``, ``util.cmap``, ``model.loss``, the dataset helpers ... all stay
    reachable -- and only the hot-path symbols are replaced:

        model.make_model, model.PixelNeRFNet (and model.models.PixelNeRFNet)   src/model/__init__.py:1-11
        render.NeRFRenderer (and render.nerf.NeRFRenderer)                     src/render/__init__.py:1
        util.gen_rays (device kernel for CUDA poses, the reference's own code otherwise)   src/util/util.py:238-276

    Without a reference tree the mirrors themselves are registered as ``model`` / ``render`` / ``util`` so that
    code written against those names still imports.  Returns the three modules now registered."""
    if reference_src is not None and reference_src not in sys.path:
        sys.path.insert(0, reference_src)
    _third_party_shims()
    from . import model as our_model, render as our_render, util as our_util
    ref_util = _import_reference("util")
    ref_model = _import_reference("model") if ref_util is not None else None
    ref_render = _import_reference("render") if ref_util is not None else None
    if ref_util is None or ref_model is None or ref_render is None:
        sys.modules["model"], sys.modules["render"], sys.modules["util"] = our_model, our_render, our_util
        return our_model, our_render, our_util

    ref_model.make_model = our_model.make_model
    ref_model.PixelNeRFNet = our_model.PixelNeRFNet
    if hasattr(ref_model, "models"):
        ref_model.models.PixelNeRFNet = our_model.PixelNeRFNet
    ref_render.NeRFRenderer = our_render.NeRFRenderer
    if hasattr(ref_render, "nerf"):
        ref_render.nerf.NeRFRenderer = our_render.NeRFRenderer

    ref_gen_rays = getattr(ref_util.gen_rays, "_pixelnerf_b200_wrapped", ref_util.gen_rays)

    def gen_rays(poses, width, height, focal, z_near, z_far, c=None, ndc=False):
        if poses.is_cuda and not ndc:
            return our_util.gen_rays(poses, width, height, focal, z_near, z_far, c=c)
        return ref_gen_rays(poses, width, height, focal, z_near, z_far, c=c, ndc=ndc)

    gen_rays.__doc__ = ref_gen_rays.__doc__
    gen_rays._pixelnerf_b200_wrapped = ref_gen_rays
    ref_util.gen_rays = gen_rays
    if hasattr(ref_util, "util"):
        ref_util.util.gen_rays = gen_rays
    return ref_model, ref_render, ref_util
