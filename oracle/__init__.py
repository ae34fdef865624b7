"""CPU oracle for the pixelNeRF render hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``pixelnerf_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
the timed CPU baseline.

The reference (etiiiR/pixel-nerf) is pure Python/PyTorch and has no tests or
golden vectors of its own, so parity is pinned by golden
fixtures generated from the live reference in the authoring container
(``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``); the oracle is
checked against those fixtures in ``tests/test_oracle_golden.py``.
"""
from .render import (  # noqa: F401
    sample_coarse,
    sample_fine,
    sample_fine_depth,
    merge_sorted,
    composite,
    render_rays,
)
from .model import (  # noqa: F401
    camera_state,
    point_features,
    positional_encoding,
    resnetfc_forward,
    net_forward,
)
from .sampling_np import (  # noqa: F401
    aten_row_sum_f32,
    sample_coarse_np,
    sample_fine_np,
    sample_fine_depth_np,
)
