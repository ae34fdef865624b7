"""pixelnerf_b200 -- B200 (sm_100a) implementation of pixelNeRF's render hot path.

Drop-in use from the reference's scripts (which do ``sys.path.insert(0, "src")`` and then
``from model import make_model; from render import NeRFRenderer; import util``):

    import pixelnerf_b200
    pixelnerf_b200.install()      # registers model / render / util / pyhocon-compatible conf

See INTEGRATION.md.
"""
import sys

__version__ = "0.1.0"


def install():
    """Make ``import model``, ``import render`` and ``import util`` resolve to this package."""
    from . import model, render, util, conf
    sys.modules["model"] = model
    sys.modules["render"] = render
    sys.modules["util"] = util
    if "pyhocon" not in sys.modules:
        sys.modules["pyhocon"] = conf
    return model, render, util
