"""Video rendering helper: the loop of eval/gen_video.py:166-237 with both ends on the device.

``render_video`` builds the rays of every target pose with the ray-generation kernel (``util.gen_rays``,
src/util/util.py:238-276), renders them in ``ray_batch_size`` chunks through a bound renderer
(``bind_parallel(net, gpus, simple_output=True)``, eval/gen_video.py:213-217) and converts the frames with
``pn_frame_to_u8`` (``(frames * 255).astype(np.uint8)``, eval/gen_video.py:234-236) before the device-to-host copy,
which therefore moves 3 bytes per pixel instead of 12.
"""
import torch

from . import _lib
from .util import util as _util


def render_video(render_par, poses, width, height, focal, z_near, z_far, c=None, ray_batch_size=50000,
                 device=None, return_depth=False):
    """poses (NV, 4, 4) camera-to-world -> uint8 frames (NV, H, W, 3) on the host (and float depth (NV, H, W) on the
    device when asked).  ``render_par`` must have been bound with ``simple_output=True``; call ``net.encode`` first."""
    if device is None:
        device = next((b.device for b in render_par.buffers()), torch.device("cuda"))
    poses = poses.to(device=device, dtype=torch.float32)
    renderer, net = getattr(render_par, "renderer", None), getattr(render_par, "net", None)
    fused = (renderer is not None and net is not None and getattr(net, "mlp_precision", "fp32") == "fp16"
             and getattr(net, "_cam", None) is not None and net._cam["n_objects"] == 1 and not renderer.training)
    if fused:
        # one GPU, tensor-core path: the rays are generated inside the render kernels (NeRFRenderer.render_frames),
        # a frame costs 64 bytes of input instead of 32 bytes per ray
        out = renderer.render_frames(net, poses, width, height, focal, z_near, z_far, c=c)
        part = out.fine if renderer.using_fine else out.coarse
        rgb_all, depth_all = part.rgb.reshape(-1, 3), part.depth.reshape(-1)
    else:
        rays = _util.gen_rays(poses, width, height, focal, z_near, z_far, c=c).reshape(-1, 8)
        rgb_parts, depth_parts = [], []
        with torch.no_grad():
            for chunk in torch.split(rays, ray_batch_size, dim=0):
                rgb, depth = render_par(chunk[None])
                rgb_parts.append(rgb[0])
                if return_depth:
                    depth_parts.append(depth[0])
        rgb_all = torch.cat(rgb_parts)
        depth_all = torch.cat(depth_parts) if return_depth else None
    frames = _lib.frame_to_u8(rgb_all).view(-1, height, width, 3)
    host = torch.empty(frames.shape, dtype=torch.uint8, pin_memory=True)
    host.copy_(frames, non_blocking=True)
    torch.cuda.current_stream(frames.device).synchronize()
    if return_depth:
        return host, depth_all.view(-1, height, width)
    return host
