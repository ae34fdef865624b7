"""Profiler ranges with the reference's names.

The reference marks its hot path with ``torch.autograd.profiler.record_function`` (src/render/nerf.py:175,264,
src/model/models.py:156, src/model/encoder.py:90,94, src/model/code.py:36, src/model/resnetfc.py:54,139).  The same
names are emitted here -- as ``record_function`` ranges (torch profiler) and as NVTX ranges (nsys / ncu
``--nvtx``) -- around the calls that replace those functions; inside the one-call render the C side pushes the same
NVTX names around its launches (csrc/pn_render.cu).
"""
import contextlib

import torch
from torch.autograd import profiler


@contextlib.contextmanager
def prof_range(name):
    nvtx = torch.cuda.is_available()
    if nvtx:
        torch.cuda.nvtx.range_push(name)
    try:
        with profiler.record_function(name):
            yield
    finally:
        if nvtx:
            torch.cuda.nvtx.range_pop()
