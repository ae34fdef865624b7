"""ctypes binding of libpixelnerf_b200.so (the C ABI in include/pixelnerf_b200.h).

There is deliberately no fallback: if the shared library is missing or a tensor is
not on a CUDA device the call raises.  torch is used only to own device memory and
to supply the current stream.
"""
import collections
import ctypes as C
import functools
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpixelnerf_b200.so")

DTYPE_F32, DTYPE_F16 = 0, 1
D_HIDDEN, D_LATENT, D_CODE, N_BLOCKS, COMBINE_LAYER = 512, 512, 42, 5, 3

_vp, _i64, _int, _u64, _f32, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_uint64, C.c_float, C.c_size_t


class Camera(C.Structure):
    _fields_ = [("poses", _vp), ("focal", _vp), ("c", _vp), ("scale_x", _f32), ("scale_y", _f32),
                ("hl", _int), ("wl", _int), ("n_objects", _int), ("n_views", _int)]


class MlpTcDebug(C.Structure):
    _fields_ = [("h", _vp), ("stage", _int)]


class MlpWeights(C.Structure):
    _fields_ = [("lin_in_w", _vp), ("lin_in_b", _vp),
                ("fc0_w", _vp * N_BLOCKS), ("fc0_b", _vp * N_BLOCKS),
                ("fc1_w", _vp * N_BLOCKS), ("fc1_b", _vp * N_BLOCKS),
                ("lin_z_w", _vp * COMBINE_LAYER), ("lin_z_b", _vp * COMBINE_LAYER),
                ("lin_out_w", _vp), ("lin_out_b", _vp)]


class MlpGrads(C.Structure):
    _fields_ = MlpWeights._fields_


class RenderConfig(C.Structure):
    _fields_ = [("n_coarse", _int), ("n_imp", _int), ("n_depth", _int), ("depth_std", _f32), ("lindisp", _int),
                ("white_bkgd", _int), ("seed", _u64)]


class RenderDraws(C.Structure):
    _fields_ = [("u_coarse", _vp), ("u_bin", _vp), ("u_in", _vp), ("noise", _vp)]


class RenderMlp(C.Structure):
    _fields_ = [("packed", _vp), ("tables", _vp)]


class RenderOut(C.Structure):
    _fields_ = [("rgb", _vp), ("depth", _vp), ("weights", _vp), ("z", _vp)]


class Frames(C.Structure):
    _fields_ = [("poses", _vp), ("pose_stride", _int), ("n_frames", _int), ("width", _int), ("height", _int),
                ("fx", _f32), ("fy", _f32), ("cx", _f32), ("cy", _f32), ("z_near", _f32), ("z_far", _f32)]


class RenderEvents(C.Structure):
    _fields_ = [("taps", (_vp * 2) * 2), ("mlp", (_vp * 2) * 2), ("composite", (_vp * 2) * 2),
                ("sample", (_vp * 2) * 2)]


_SIGNATURES = {
    "pn_version": (_int, []),
    "pn_last_error": (C.c_char_p, []),
    "pn_launch_count": (_u64, []),
    "pn_set_device_seed": (None, [_vp]),
    "pn_sample_coarse": (_int, [_vp, _vp, _vp, _vp, _i64, _int, _int, _u64, _vp]),
    "pn_sample_fine": (_int, [_vp] * 10 + [_i64, _int, _int, _int, _f32, _int, _u64, _vp]),
    "pn_prepare_latent": (_int, [_vp, _vp, _int, _int, _int, _int, _int, _vp]),
    "pn_point_features": (_int, [C.POINTER(Camera), _vp, _int, _vp, _vp, _int, _vp, _vp, _i64, _vp, _vp,
                                 _int, _int, _vp]),
    "pn_mlp_fp32_workspace_bytes": (_sz, [_i64]),
    "pn_mlp_forward_fp32": (_int, [C.POINTER(MlpWeights), _vp, _vp, _int, _int, _int, _i64, _vp, _vp, _vp,
                                   _sz, _vp]),
    "pn_point_features_bwd": (_int, [C.POINTER(Camera), _vp, _vp, _int, _vp, _i64, _vp, _vp, _vp]),
    "pn_depth_sample_bwd": (_int, [C.POINTER(Camera), _vp, _vp, _vp, _int, _vp, _int, _vp, _vp, _int, _vp, _vp, _i64,
                                   _vp, _vp]),
    "pn_mlp_backward_gh0_offset": (_sz, [_i64, _int]),
    "pn_mlp_backward_tc_gh0_offset": (_sz, [_i64, _int]),
    "pn_mlp_train_acts_bytes": (_sz, [_i64]),
    "pn_mlp_forward_fp32_train": (_int, [C.POINTER(MlpWeights), _vp, _vp, _int, _int, _int, _i64, _vp, _vp, _vp,
                                         _sz, _vp]),
    "pn_mlp_backward_workspace_bytes": (_sz, [_i64]),
    "pn_mlp_backward_fp32": (_int, [C.POINTER(MlpWeights), C.POINTER(MlpWeights), _vp, _vp, _int, _int, _int, _i64,
                                    _vp, _vp, _vp, C.POINTER(MlpGrads), _vp, _vp, _sz, _vp]),
    "pn_mlp_tc_packed_bytes": (_sz, []),
    "pn_prepare_weights": (_int, [C.POINTER(MlpWeights), _vp, _sz, _vp]),
    "pn_mlp_tc_workspace_bytes": (_sz, [_i64, _int]),
    "pn_mlp_forward_tc": (_int, [_vp, _vp, _i64, _vp, _int, _int, _i64, _vp, _vp, _vp, _sz, _vp, _vp]),
    # no-allocation variant of pn_mlp_forward_tc (scratch in a static device array, csrc/pn_noalloc.cu)
    "pn_static_scratch_bytes": (_sz, []),
    "pn_mlp_forward_tc_static": (_int, [_vp, _vp, _i64, _vp, _int, _int, _i64, _vp, _vp, _vp]),
    # precise tensor-core chain (mlp_precision = fp16x3)
    "pn_prepare_latent_tables_f32": (_int, [C.POINTER(MlpWeights), _vp, _i64, _vp, _vp]),
    "pn_point_taps_f32": (_int, [C.POINTER(Camera), _vp, _vp, _int, _vp, _vp, _i64, _vp, _vp, _int, _vp]),
    "pn_mlp_tc_packed_x3_bytes": (_sz, []),
    "pn_prepare_weights_x3": (_int, [C.POINTER(MlpWeights), _vp, _sz, _vp]),
    "pn_mlp_forward_tc_x3": (_int, [_vp, _vp, _i64, _vp, _vp, _int, _int, _i64, _vp, _vp, _vp, _sz, _vp]),
    "pn_mlp_forward_tc_x2": (_int, [_vp, _vp, _i64, _vp, _vp, _int, _int, _i64, _vp, _vp, _vp, _sz, _vp]),
    # n_views > 1 in one launch: grouped row order + on-chip view mean
    "pn_point_taps_grouped_rows": (_i64, [_i64, _int]),
    "pn_point_taps_grouped": (_int, [C.POINTER(Camera), _vp, _vp, _int, _vp, _vp, _i64, _vp, _vp]),
    "pn_mlp_tc_grouped_workspace_bytes": (_sz, []),
    "pn_mlp_forward_tc_grouped": (_int, [_vp, _vp, _i64, _vp, _int, _int, _i64, _vp, _vp, _vp, _sz, _vp]),
    "pn_point_taps": (_int, [C.POINTER(Camera), _vp, _vp, _int, _vp, _vp, _i64, _vp, _vp]),
    "pn_latent_tables_bytes": (_sz, [_i64]),
    "pn_prepare_latent_tables": (_int, [C.POINTER(MlpWeights), _vp, _i64, _vp, _vp, _sz, _vp]),
    "pn_latent_tables_tc_workspace_bytes": (_sz, [_i64]),
    "pn_prepare_latent_tables_tc": (_int, [C.POINTER(MlpWeights), _vp, _i64, _vp, _vp, _sz, _vp]),
    "pn_composite_fwd": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _int, _int, _vp]),
    "pn_composite_bwd": (_int, [_vp] * 8 + [_i64, _int, _int, _vp]),
    "pn_mlp_tc_acts_bytes": (_sz, [_i64]),
    "pn_mlp_forward_tc_train": (_int, [_vp, _vp, _i64, _vp, _int, _int, _i64, _vp, _vp, _vp, _sz, _vp, _sz, _vp]),
    "pn_mlp_tc_w16_bytes": (_sz, []),
    "pn_prepare_weights_bf16": (_int, [C.POINTER(MlpWeights), _vp, _sz, _vp]),
    "pn_mlp_backward_tc_workspace_bytes": (_sz, [_i64, _int]),
    "pn_mlp_backward_tc": (_int, [C.POINTER(MlpWeights), _vp, _vp, _vp, _int, _int, _i64, _vp, _vp, _vp,
                                  C.POINTER(MlpGrads), _vp, _vp, _sz, _vp]),
    "pn_gemm_tc_rows": (_int, [_vp, _int, _vp, _int, _i64, _int, _int, _vp, _vp, _int, _vp, _vp, _f32, _vp, _vp, _vp,
                               _vp]),
    "pn_gemm_tc_rows_slabs": (_int, [_vp, _i64, _int, _vp, _i64, _int, _int, _vp, _int, _f32, _vp, _vp]),
    "pn_gemm_tc_wgrad": (_int, [_vp, _int, _int, _vp, _int, _int, _i64, _int, _vp, _int, _vp, _vp]),
    "pn_gen_rays": (_int, [_vp, _int, _int, _int, _int, _f32, _f32, _f32, _f32, _f32, _f32, _vp, _vp]),
    "pn_frame_to_u8": (_int, [_vp, _vp, _i64, _vp]),
    "pn_encoder_latent": (_int, [_vp, _vp, _vp, _vp, _int, _int, _vp, _int, _vp]),
    "pn_render_rays_workspace_bytes": (_sz, [_i64, _int, C.POINTER(RenderConfig)]),
    "pn_render_frames": (_int, [C.POINTER(Camera), C.POINTER(RenderMlp), C.POINTER(RenderMlp), _i64, C.POINTER(Frames),
                                _i64, _i64, _vp, C.POINTER(RenderConfig), C.POINTER(RenderDraws), C.POINTER(RenderOut),
                                C.POINTER(RenderOut), _vp, _sz, C.POINTER(RenderEvents), _vp]),
    "pn_render_rays": (_int, [C.POINTER(Camera), C.POINTER(RenderMlp), C.POINTER(RenderMlp), _i64, _vp, _i64, _vp,
                              C.POINTER(RenderConfig), C.POINTER(RenderDraws), C.POINTER(RenderOut),
                              C.POINTER(RenderOut), _vp, _sz, C.POINTER(RenderEvents), _vp]),
}

_lib = None

# bench.py sets PROFILE = {} to collect CUDA-event pairs per kernel family (recorded on the
# launching stream, read after the step's final synchronise); None = no events.
PROFILE = None


class _Timed:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if PROFILE is not None:
            self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            self.ev[0].record()

    def __exit__(self, *exc):
        if PROFILE is not None:
            self.ev[1].record()
            PROFILE.setdefault(self.name, []).append(self.ev)


def exported_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Load the shared library once; raise (never fall back) if it is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C pixelnerf_b200/csrc).  pixelnerf_b200 has no non-CUDA fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def launch_count():
    return int(lib().pn_launch_count())


def set_device_seed(word):
    """Sampling launches made by this host thread from now on add the int64 device word to their host seed
    (include/pixelnerf_b200.h: pn_set_device_seed); None restores the host seed alone."""
    if word is not None and (word.dtype != torch.int64 or word.numel() != 1 or not word.is_cuda):
        raise RuntimeError("device seed must be a one-element int64 CUDA tensor")
    lib().pn_set_device_seed(None if word is None else word.data_ptr())


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib().pn_last_error().decode()}")


def _ptr(t, dtype=torch.float32):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("pixelnerf_b200 kernels need CUDA tensors (there is no CPU path)")
    if t.dtype != dtype or not t.is_contiguous():
        raise RuntimeError(f"expected contiguous {dtype} tensor, got {t.dtype} contiguous={t.is_contiguous()}")
    return t.data_ptr()


def _f32c(t):
    """Contiguous fp32 view of `t` without autograd history (no copy, no dispatcher call, when it already is one)."""
    if t is None:
        return None
    if t.dtype == torch.float32 and t.is_contiguous() and not t.requires_grad:
        return t
    return t.detach().to(torch.float32).contiguous()


def _stream():
    # the current stream of the CURRENT device: every launching entry point below runs under _on_device, which
    # makes the tensors' device current first
    return torch.cuda.current_stream().cuda_stream


def _tensor_device(objs):
    """The one CUDA device every tensor among `objs` (nested dicts / lists / tuples included) lives on; raises if
    they disagree (a kernel launched on one GPU's stream with another GPU's pointers faults or, with peer access
    enabled, silently reads across NVLink)."""
    dev = None
    stack = list(objs)
    while stack:
        o = stack.pop()
        if torch.is_tensor(o):
            if o.is_cuda:
                if dev is None:
                    dev = o.device
                elif o.device != dev:
                    raise RuntimeError(f"pixelnerf_b200: tensors of one call live on {dev} and {o.device}")
        elif isinstance(o, dict):
            stack.extend(o.values())
        elif isinstance(o, (list, tuple)):
            stack.extend(o)
    return dev


def _on_device(fn):
    """Run `fn` with its tensors' device current: the reference's callers pick a device with
    ``util.get_cuda(args.gpu_id[0])`` and never call ``torch.cuda.set_device`` (eval/gen_video.py:66), and the
    in-process multi-GPU wrapper drives several devices from one process."""
    @functools.wraps(fn)
    def call(*args, **kwargs):
        dev = _tensor_device(args + tuple(kwargs.values()))
        if dev is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return call


_LIN_CACHE = {}


def _linspace(n_coarse, device):
    key = (n_coarse, str(device))
    if key not in _LIN_CACHE:  # host-made table, see SURVEY 8a(a2): never re-derive torch.linspace
        _LIN_CACHE[key] = torch.linspace(0, 1 - 1.0 / n_coarse, n_coarse).to(device)
    return _LIN_CACHE[key]


def sample_coarse(rays, n_coarse, lindisp=False, u=None, seed=0):
    rays, u = _f32c(rays), _f32c(u)
    z = torch.empty(rays.shape[0], n_coarse, device=rays.device, dtype=torch.float32)
    _check(lib().pn_sample_coarse(_ptr(rays), _ptr(u), _ptr(_linspace(n_coarse, rays.device)), _ptr(z),
                                  rays.shape[0], n_coarse, int(bool(lindisp)), seed, _stream()), "pn_sample_coarse")
    return z


def sample_fine(rays, z_coarse, weights, depth, n_imp, n_depth, depth_std, lindisp=False,
                u_bin=None, u_in=None, noise=None, seed=0, want_parts=False):
    rays, z_coarse, weights, depth = _f32c(rays), _f32c(z_coarse), _f32c(weights), _f32c(depth)
    u_bin, u_in, noise = _f32c(u_bin), _f32c(u_in), _f32c(noise)
    r, kc = z_coarse.shape
    dev = rays.device
    z_sorted = torch.empty(r, kc + n_imp + n_depth, device=dev, dtype=torch.float32)
    z_imp = torch.empty(r, n_imp, device=dev, dtype=torch.float32) if want_parts else None
    z_dep = torch.empty(r, n_depth, device=dev, dtype=torch.float32) if want_parts else None
    _check(lib().pn_sample_fine(_ptr(rays), _ptr(z_coarse), _ptr(weights), _ptr(depth), _ptr(u_bin), _ptr(u_in),
                                _ptr(noise), _ptr(z_sorted), _ptr(z_imp), _ptr(z_dep), r, kc, n_imp, n_depth,
                                float(depth_std), int(bool(lindisp)), seed, _stream()), "pn_sample_fine")
    return (z_sorted, z_imp, z_dep) if want_parts else z_sorted


def prepare_latent(latent_nchw, dtype=torch.float32):
    lat = _f32c(latent_nchw)
    v, ch, hl, wl = lat.shape
    if ch != D_LATENT:
        raise NotImplementedError(f"latent size {ch} != {D_LATENT}: only the 4-level ResNet34 latent is supported")
    out = torch.empty(v, hl, wl, ch, device=lat.device, dtype=dtype)
    code = DTYPE_F32 if dtype == torch.float32 else DTYPE_F16
    _check(lib().pn_prepare_latent(_ptr(lat), out.data_ptr(), v, ch, hl, wl, code, _stream()), "pn_prepare_latent")
    return out


def point_features(cam_bufs, latent_nhwc, ppo, *, rays=None, z=None, xyz=None, viewdirs=None, ldx=44):
    """cam_bufs: dict(poses (V,3,4), focal (V,2), c (V,2), scale_x, scale_y, n_objects, n_views)."""
    dt = latent_nhwc.dtype
    code = DTYPE_F32 if dt == torch.float32 else DTYPE_F16
    v, hl, wl, _ = latent_nhwc.shape
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    if v != cam.n_objects * cam.n_views:
        raise RuntimeError(f"latent holds {v} maps but camera state has {cam.n_objects}x{cam.n_views} views")
    n_rows = v * ppo
    dev = latent_nhwc.device
    feat_z = torch.empty(n_rows, D_LATENT, device=dev, dtype=dt)
    feat_x = torch.empty(n_rows, ldx, device=dev, dtype=dt)
    rays, z, xyz, viewdirs = _f32c(rays), _f32c(z), _f32c(xyz), _f32c(viewdirs)
    k = 0 if z is None else z.shape[-1]
    with _Timed("features"):
        _check(lib().pn_point_features(C.byref(cam), latent_nhwc.data_ptr(), code, _ptr(rays), _ptr(z), k,
                                       _ptr(xyz), _ptr(viewdirs), ppo, feat_z.data_ptr(), feat_x.data_ptr(), ldx,
                                       code, _stream()), "pn_point_features")
    return feat_z, feat_x


_PACK32_CACHE = collections.OrderedDict()


def pack_weights_fp32(w):
    """dict of un-prefixed ResnetFC tensors (CUDA fp32 contiguous) -> (MlpWeights, keepalive).  The last few results
    are kept, keyed on every tensor's (data_ptr, _version): a training step asks for the same struct eight times
    (tables, forward, backward of both MLPs) and building it is ~100 torch calls."""
    key = tuple((v.data_ptr(), v._version, v.dtype) for v in w.values())
    hit = _PACK32_CACHE.get(key)
    if hit is not None:
        _PACK32_CACHE.move_to_end(key)
        return hit
    hit = _pack_weights_fp32(w)
    _PACK32_CACHE[key] = hit       # the keepalive dict pins the storages, so a cached data_ptr cannot be recycled
    while len(_PACK32_CACHE) > 8:
        _PACK32_CACHE.popitem(last=False)
    return hit


def _pack_weights_fp32(w):
    keep = {k: _f32c(v) for k, v in w.items()}
    s = MlpWeights()
    s.lin_in_w, s.lin_in_b = _ptr(keep["lin_in.weight"]), _ptr(keep["lin_in.bias"])
    for b in range(N_BLOCKS):
        s.fc0_w[b], s.fc0_b[b] = _ptr(keep[f"blocks.{b}.fc_0.weight"]), _ptr(keep[f"blocks.{b}.fc_0.bias"])
        s.fc1_w[b], s.fc1_b[b] = _ptr(keep[f"blocks.{b}.fc_1.weight"]), _ptr(keep[f"blocks.{b}.fc_1.bias"])
    for b in range(COMBINE_LAYER):
        s.lin_z_w[b], s.lin_z_b[b] = _ptr(keep[f"lin_z.{b}.weight"]), _ptr(keep[f"lin_z.{b}.bias"])
    s.lin_out_w, s.lin_out_b = _ptr(keep["lin_out.weight"]), _ptr(keep["lin_out.bias"])
    return s, keep


def mlp_forward_fp32(packed, feat_z, feat_x, n_objects, n_views, ppo, want_pre=False):
    s, _keep = packed
    n_rows = n_objects * n_views * ppo
    dev = feat_z.device
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    pre = torch.empty_like(out) if want_pre else None
    nbytes = lib().pn_mlp_fp32_workspace_bytes(n_rows)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    with _Timed("mlp"):
        _check(lib().pn_mlp_forward_fp32(C.byref(s), _ptr(feat_z), _ptr(feat_x), feat_x.shape[1], n_objects,
                                         n_views, ppo, _ptr(out), _ptr(pre), ws.data_ptr(), nbytes, _stream()),
               "pn_mlp_forward_fp32")
    return (out, pre) if want_pre else out


_WEIGHT_NAMES = (["lin_in.weight", "lin_in.bias"]
                 + [f"blocks.{b}.fc_{i}.{k}" for b in range(N_BLOCKS) for i in (0, 1) for k in ("weight", "bias")]
                 + [f"lin_z.{b}.{k}" for b in range(COMBINE_LAYER) for k in ("weight", "bias")]
                 + ["lin_out.weight", "lin_out.bias"])


def weight_names():
    """Un-prefixed ResnetFC state_dict names, the order used for flat parameter / gradient lists."""
    return list(_WEIGHT_NAMES)


def _fill_struct(s, t):
    s.lin_in_w, s.lin_in_b = _ptr(t["lin_in.weight"]), _ptr(t["lin_in.bias"])
    for b in range(N_BLOCKS):
        s.fc0_w[b], s.fc0_b[b] = _ptr(t[f"blocks.{b}.fc_0.weight"]), _ptr(t[f"blocks.{b}.fc_0.bias"])
        s.fc1_w[b], s.fc1_b[b] = _ptr(t[f"blocks.{b}.fc_1.weight"]), _ptr(t[f"blocks.{b}.fc_1.bias"])
    for b in range(COMBINE_LAYER):
        s.lin_z_w[b], s.lin_z_b[b] = _ptr(t[f"lin_z.{b}.weight"]), _ptr(t[f"lin_z.{b}.bias"])
    s.lin_out_w, s.lin_out_b = _ptr(t["lin_out.weight"]), _ptr(t["lin_out.bias"])
    return s


def mlp_forward_fp32_train(packed, feat_z, feat_x, n_objects, n_views, ppo):
    """fp32 forward that keeps the activations: returns (out (P,4), acts buffer)."""
    s, _keep = packed
    n_rows = n_objects * n_views * ppo
    dev = feat_z.device
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    nbytes = lib().pn_mlp_train_acts_bytes(n_rows)
    acts = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    _check(lib().pn_mlp_forward_fp32_train(C.byref(s), _ptr(feat_z), _ptr(feat_x), feat_x.shape[1], n_objects, n_views,
                                           ppo, _ptr(out), None, acts.data_ptr(), nbytes, _stream()),
           "pn_mlp_forward_fp32_train")
    return out, acts


def mlp_backward_fp32(packed, weights, feat_z, feat_x, n_objects, n_views, ppo, acts, out, g_out, grads,
                      want_g_feat_z, want_g_h0=False):
    """weights / grads: dicts keyed by weight_names(); grads are accumulated into (+=).  Returns dL/d feat_z
    (or None); with want_g_h0 also dL/dh_0 (rows, 512) fp32, a view into the call's workspace."""
    s, _keep = packed
    wt = {k: (v.detach().t().contiguous() if v.dim() == 2 and v.shape[0] == v.shape[1] else v.detach().contiguous())
          for k, v in weights.items()}
    st = _fill_struct(MlpWeights(), wt)
    sg = _fill_struct(MlpGrads(), grads)
    n_rows = n_objects * n_views * ppo
    dev = feat_z.device
    g_fz = torch.empty(n_rows, D_LATENT, device=dev, dtype=torch.float32) if want_g_feat_z else None
    nbytes = lib().pn_mlp_backward_workspace_bytes(n_rows)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    _check(lib().pn_mlp_backward_fp32(C.byref(s), C.byref(st), _ptr(feat_z), _ptr(feat_x), feat_x.shape[1], n_objects,
                                      n_views, ppo, acts.data_ptr(), _ptr(out), _ptr(_f32c(g_out)), C.byref(sg),
                                      _ptr(g_fz), ws.data_ptr(), nbytes, _stream()), "pn_mlp_backward_fp32")
    if want_g_h0:
        off = lib().pn_mlp_backward_gh0_offset(n_rows, n_views)
        return g_fz, ws[off:off + n_rows * D_HIDDEN * 4].view(torch.float32).view(n_rows, D_HIDDEN)
    return g_fz


def point_features_bwd(cam_bufs, g_latent_nhwc, ppo, g_feat_z, *, rays=None, z=None, xyz=None):
    v, hl, wl, _ = g_latent_nhwc.shape
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    rays, z, xyz = _f32c(rays), _f32c(z), _f32c(xyz)
    k = 0 if z is None else z.shape[-1]
    _check(lib().pn_point_features_bwd(C.byref(cam), _ptr(rays), _ptr(z), k, _ptr(xyz), ppo, _ptr(g_feat_z),
                                       _ptr(g_latent_nhwc), _stream()), "pn_point_features_bwd")


def depth_sample_bwd(cam_bufs, latent_nhwc, rays, z_sorted, z_depth, g_z, g_h0, g_feat_z, lin_in_w, g_depth):
    """g_depth (R) += dL/d(coarse depth) through the depth-guided samples of the fine pass (nerf.py:150-161)."""
    v, hl, wl, _ = latent_nhwc.shape
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    rays, z_sorted, z_depth, g_z = _f32c(rays), _f32c(z_sorted), _f32c(z_depth), _f32c(g_z)
    r, k = z_sorted.shape
    if g_h0.dtype not in (torch.float32, torch.bfloat16) or not g_h0.is_contiguous():
        raise RuntimeError("depth_sample_bwd: g_h0 must be contiguous fp32 or bf16")
    if g_depth.dtype != torch.float32 or g_depth.shape != (r,) or not g_depth.is_contiguous():
        raise RuntimeError("depth_sample_bwd: g_depth must be a contiguous fp32 (R,) tensor")
    fmt = 0 if g_h0.dtype == torch.float32 else 2
    _check(lib().pn_depth_sample_bwd(C.byref(cam), _ptr(latent_nhwc), _ptr(rays), _ptr(z_sorted), k, _ptr(z_depth),
                                     z_depth.shape[1], _ptr(g_z), g_h0.data_ptr(), fmt, _ptr(g_feat_z),
                                     _ptr(_f32c(lin_in_w)), r // cam_bufs["n_objects"], _ptr(g_depth), _stream()),
           "pn_depth_sample_bwd")


def point_taps(cam_bufs, hl, wl, ppo, *, rays=None, z=None, xyz=None, viewdirs=None):
    """Tensor-core path: feat_x (N,64) fp16 = positional code in columns [0,42) and, in bytes [96,128) of each
    row, the taps record (4 int32 texel indices + 4 fp32 bilinear weights)."""
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    n_rows = cam.n_objects * cam.n_views * ppo
    dev = cam_bufs["poses"].device
    feat_x = torch.empty(n_rows, 64, device=dev, dtype=torch.float16)
    rays, z, xyz, viewdirs = _f32c(rays), _f32c(z), _f32c(xyz), _f32c(viewdirs)
    k = 0 if z is None else z.shape[-1]
    with _Timed("features"):
        _check(lib().pn_point_taps(C.byref(cam), _ptr(rays), _ptr(z), k, _ptr(xyz), _ptr(viewdirs), ppo,
                                   feat_x.data_ptr(), _stream()), "pn_point_taps")
    return feat_x


def point_taps_grouped(cam_bufs, hl, wl, ppo, *, rays=None, z=None, xyz=None, viewdirs=None):
    """point_taps in GROUPED row order for mlp_forward_tc_grouped (n_views > 1): row (S * n_views + v) * 128 + m =
    view v of point 128 S + m of the flattened (object, point) index (include/pixelnerf_b200.h)."""
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    n_rows = int(lib().pn_point_taps_grouped_rows(cam.n_objects * ppo, cam.n_views))
    feat_x = torch.empty(n_rows, 64, device=cam_bufs["poses"].device, dtype=torch.float16)
    rays, z, xyz, viewdirs = _f32c(rays), _f32c(z), _f32c(xyz), _f32c(viewdirs)
    k = 0 if z is None else z.shape[-1]
    with _Timed("features"):
        _check(lib().pn_point_taps_grouped(C.byref(cam), _ptr(rays), _ptr(z), k, _ptr(xyz), _ptr(viewdirs), ppo,
                                           feat_x.data_ptr(), _stream()), "pn_point_taps_grouped")
    return feat_x


def mlp_forward_tc_grouped(packed, tables, feat_x_grouped, n_objects, n_views, ppo, want_pre=False):
    """The tensor-core chain for n_views > 1 in one launch (view mean on chip); feat_x_grouped from
    point_taps_grouped.  -> (SB*ppo, 4) [sigmoid rgb, relu sigma]."""
    dev = feat_x_grouped.device
    n_rows = int(lib().pn_point_taps_grouped_rows(n_objects * ppo, n_views))
    if feat_x_grouped.shape != (n_rows, 64) or feat_x_grouped.dtype != torch.float16 or not feat_x_grouped.is_contiguous():
        raise RuntimeError("grouped tensor-core MLP needs contiguous fp16 feat_x of shape (grouped rows, 64)")
    if tables.dtype != torch.float16 or tables.dim() != 3 or tables.shape[0] != COMBINE_LAYER or not tables.is_contiguous():
        raise RuntimeError("tensor-core MLP needs contiguous fp16 lin_z tables of shape (3, texels, 512)")
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    pre = torch.empty_like(out) if want_pre else None
    nbytes = lib().pn_mlp_tc_grouped_workspace_bytes()
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    with _Timed("mlp"):
        _check(lib().pn_mlp_forward_tc_grouped(packed.data_ptr(), tables.data_ptr(), tables.shape[1],
                                               _ptr(feat_x_grouped, torch.float16), n_objects, n_views, ppo, _ptr(out),
                                               _ptr(pre), ws.data_ptr(), nbytes, _stream()), "pn_mlp_forward_tc_grouped")
    return (out, pre) if want_pre else out


def prepare_latent_tables_f32(w, latent_nhwc):
    """Per-texel lin_z tables (3, V*Hl*Wl, 512) in fp32 for the precise chain (mlp_precision = fp16x3)."""
    s, keep = pack_weights_fp32(w)
    if latent_nhwc.dtype != torch.float32:
        raise RuntimeError("prepare_latent_tables_f32 needs the fp32 channels-last latent")
    v, hl, wl, ch = latent_nhwc.shape
    n_tex = v * hl * wl
    tables = torch.empty(COMBINE_LAYER, n_tex, D_HIDDEN, device=latent_nhwc.device, dtype=torch.float32)
    _check(lib().pn_prepare_latent_tables_f32(C.byref(s), _ptr(latent_nhwc), n_tex, _ptr(tables), _stream()),
           "pn_prepare_latent_tables_f32")
    return tables


def pack_weights_x3(w):
    """Packed buffer of the precise chain: fp16 hi + lo layers, bias tables, lin_in.weight^T in fp32."""
    s, keep = pack_weights_fp32(w)
    nbytes = lib().pn_mlp_tc_packed_x3_bytes()
    packed = torch.empty(nbytes, device=keep["lin_in.weight"].device, dtype=torch.uint8)
    _check(lib().pn_prepare_weights_x3(C.byref(s), packed.data_ptr(), nbytes, _stream()), "pn_prepare_weights_x3")
    return packed


def point_taps_f32(cam_bufs, hl, wl, ppo, *, rays=None, z=None, xyz=None, viewdirs=None):
    """point_taps (n_views == 1) / point_taps_grouped (n_views > 1) plus the fp16 remainder of the positional code:
    returns (feat_x (rows, 64) fp16, feat_x_lo (rows, 64) fp16) in the same row order; columns [0, 42) of the two are
    the fp16 pair x_hi + x_lo of the unrounded code (the rest of feat_x_lo is zero)."""
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    grouped = cam.n_views > 1
    n_rows = int(lib().pn_point_taps_grouped_rows(cam.n_objects * ppo, cam.n_views)) if grouped else cam.n_objects * ppo
    dev = cam_bufs["poses"].device
    feat_x = torch.empty(n_rows, 64, device=dev, dtype=torch.float16)
    feat_x_lo = torch.empty(n_rows, 64, device=dev, dtype=torch.float16)
    rays, z, xyz, viewdirs = _f32c(rays), _f32c(z), _f32c(xyz), _f32c(viewdirs)
    k = 0 if z is None else z.shape[-1]
    with _Timed("features"):
        _check(lib().pn_point_taps_f32(C.byref(cam), _ptr(rays), _ptr(z), k, _ptr(xyz), _ptr(viewdirs), ppo,
                                       feat_x.data_ptr(), feat_x_lo.data_ptr(), int(grouped), _stream()), "pn_point_taps_f32")
    return feat_x, feat_x_lo


def mlp_forward_tc_x3(packed_x3, tables32, feat_x, feat_x_lo, n_objects, n_views, ppo, want_pre=False, passes=3):
    """The precise tensor-core chain (three fp16 hi/lo operand passes per GEMM, lin_in on the fp16 pair of the code,
    tables and head fp32); inputs from pack_weights_x3 / prepare_latent_tables_f32 / point_taps_f32.
    -> (SB*ppo, 4) [sigmoid rgb, relu sigma].
    passes=2: the two-pass form (mlp_precision = fp16x2, pn_mlp_forward_tc_x2): no a_lo pass."""
    assert passes in (2, 3)
    entry, what = (lib().pn_mlp_forward_tc_x3, "pn_mlp_forward_tc_x3") if passes == 3 else \
        (lib().pn_mlp_forward_tc_x2, "pn_mlp_forward_tc_x2")
    dev = feat_x.device
    if tables32.dtype != torch.float32 or tables32.dim() != 3 or tables32.shape[0] != COMBINE_LAYER or not tables32.is_contiguous():
        raise RuntimeError("precise tensor-core MLP needs contiguous fp32 lin_z tables of shape (3, texels, 512)")
    if feat_x.dtype != torch.float16 or feat_x.shape[1] != 64 or feat_x_lo.shape != feat_x.shape \
            or feat_x_lo.dtype != torch.float16 or not feat_x.is_contiguous() or not feat_x_lo.is_contiguous():
        raise RuntimeError("precise tensor-core MLP needs feat_x and feat_x_lo, both (rows, 64) fp16, from point_taps_f32")
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    pre = torch.empty_like(out) if want_pre else None
    nbytes = lib().pn_mlp_tc_grouped_workspace_bytes()
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    with _Timed("mlp"):
        _check(entry(packed_x3.data_ptr(), _ptr(tables32), tables32.shape[1], _ptr(feat_x, torch.float16),
                     _ptr(feat_x_lo, torch.float16),
                     n_objects, n_views, ppo, _ptr(out), _ptr(pre), ws.data_ptr(), nbytes, _stream()), what)
    return (out, pre) if want_pre else out


def prepare_latent_tables(w, latent_nhwc):
    """Per-texel lin_z tables (3, V*Hl*Wl, 512) fp16 of one MLP for the tensor-core chain.  fp32 NHWC latent: the
    fp32 CUDA-core product (once per encode: inference); fp16 NHWC latent: the tensor-core product with hi + lo
    weights (training mode, where the tables are rebuilt every step)."""
    s, keep = pack_weights_fp32(w)
    v, hl, wl, ch = latent_nhwc.shape
    n_tex = v * hl * wl
    dev = latent_nhwc.device
    tables = torch.empty(COMBINE_LAYER, n_tex, D_HIDDEN, device=dev, dtype=torch.float16)
    if latent_nhwc.dtype == torch.float16:
        nbytes = lib().pn_latent_tables_tc_workspace_bytes(n_tex)
        ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
        with _Timed("tables"):
            _check(lib().pn_prepare_latent_tables_tc(C.byref(s), _ptr(latent_nhwc, torch.float16), n_tex,
                                                     tables.data_ptr(), ws.data_ptr(), nbytes, _stream()),
                   "pn_prepare_latent_tables_tc")
        return tables
    if latent_nhwc.dtype != torch.float32:
        raise RuntimeError("prepare_latent_tables needs the fp32 or fp16 channels-last latent")
    ws = torch.empty(n_tex * D_HIDDEN, device=dev, dtype=torch.float32)
    _check(lib().pn_prepare_latent_tables(C.byref(s), latent_nhwc.data_ptr(), n_tex, tables.data_ptr(), ws.data_ptr(),
                                          ws.numel() * 4, _stream()), "pn_prepare_latent_tables")
    return tables


def pack_weights_tc(w, precision="fp16"):
    """dict of un-prefixed ResnetFC tensors -> packed fp16 layers + bias tables (one device buffer)."""
    if precision != "fp16":
        raise NotImplementedError(f"tensor-core MLP operand format {precision!r}: only fp16 is implemented")
    s, keep = pack_weights_fp32(w)
    dev = keep["lin_in.weight"].device
    nbytes = lib().pn_mlp_tc_packed_bytes()
    packed = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    _check(lib().pn_prepare_weights(C.byref(s), packed.data_ptr(), nbytes, _stream()), "pn_prepare_weights")
    return packed


def mlp_forward_tc(packed, tables, feat_x, n_objects, n_views, ppo, want_pre=False, debug_stage=None):
    """lin_z tables (3,T,512) fp16 + feat_x (N,64) fp16 with embedded taps -> (SB*ppo, 4) [sigmoid rgb, relu sigma]."""
    n_rows = n_objects * n_views * ppo
    dev = feat_x.device
    if feat_x.shape != (n_rows, 64) or feat_x.dtype != torch.float16 or not feat_x.is_contiguous():
        raise RuntimeError("tensor-core MLP needs contiguous fp16 feat_x of shape (rows, 64)")
    if tables.dtype != torch.float16 or tables.dim() != 3 or tables.shape[0] != COMBINE_LAYER or not tables.is_contiguous():
        raise RuntimeError("tensor-core MLP needs contiguous fp16 lin_z tables of shape (3, texels, 512)")
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    pre = torch.empty_like(out) if want_pre else None
    nbytes = lib().pn_mlp_tc_workspace_bytes(n_rows, n_views)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    dbg, dbg_h = None, None
    if debug_stage is not None:
        rows = n_rows if debug_stage < COMBINE_LAYER or n_views == 1 else n_objects * ppo
        dbg_h = torch.zeros(rows, D_HIDDEN, device=dev, dtype=torch.float32)
        dbg = C.byref(MlpTcDebug(dbg_h.data_ptr(), int(debug_stage)))
    if os.environ.get("PN_TC_EXPERIMENT"):   # timing experiments (wrong results), see pn_mlp_tc.cu
        dbg = C.byref(MlpTcDebug(None, {"no_epilogue": -99, "no_gather": -98, "gather_loads_only": -97, "gather_no_loads": -96}[os.environ["PN_TC_EXPERIMENT"]]))
    with _Timed("mlp"):
        _check(lib().pn_mlp_forward_tc(packed.data_ptr(), tables.data_ptr(), tables.shape[1],
                                       _ptr(feat_x, torch.float16), n_objects, n_views, ppo, _ptr(out), _ptr(pre), ws.data_ptr(), nbytes, dbg,
                                       _stream()), "pn_mlp_forward_tc")
    ret = (out, pre) if want_pre else out
    return (ret, dbg_h) if debug_stage is not None else ret


def composite_fwd(rgb_sigma, z, rays, white_bkgd):
    rgb_sigma, z, rays = _f32c(rgb_sigma), _f32c(z), _f32c(rays)
    r, k = z.shape
    dev = z.device
    weights = torch.empty(r, k, device=dev, dtype=torch.float32)
    rgb = torch.empty(r, 3, device=dev, dtype=torch.float32)
    depth = torch.empty(r, device=dev, dtype=torch.float32)
    with _Timed("composite"):
        _check(lib().pn_composite_fwd(_ptr(rgb_sigma), _ptr(z), _ptr(rays), _ptr(weights), _ptr(rgb), _ptr(depth),
                                      r, k, int(bool(white_bkgd)), _stream()), "pn_composite_fwd")
    return weights, rgb, depth


def composite_bwd(rgb_sigma, z, rays, white_bkgd, g_rgb, g_depth=None, g_weights=None, want_gz=True):
    rgb_sigma, z, rays = _f32c(rgb_sigma), _f32c(z), _f32c(rays)
    g_rgb, g_depth, g_weights = _f32c(g_rgb), _f32c(g_depth), _f32c(g_weights)
    r, k = z.shape
    g_rs = torch.empty(r, k, 4, device=z.device, dtype=torch.float32)
    g_z = torch.empty(r, k, device=z.device, dtype=torch.float32) if want_gz else None
    _check(lib().pn_composite_bwd(_ptr(rgb_sigma), _ptr(z), _ptr(rays), _ptr(g_rgb), _ptr(g_depth), _ptr(g_weights),
                                  _ptr(g_rs), _ptr(g_z), r, k, int(bool(white_bkgd)), _stream()), "pn_composite_bwd")
    return g_rs, g_z


def _new_event():
    """A timing event whose raw cudaEvent_t exists (torch creates it lazily on the first record)."""
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def frames_struct(poses, width, height, focal, z_near, z_far, c=None):
    """(Frames, keep-alive) for pn_render_frames: poses (NV, 3|4, 4) CUDA camera-to-world; focal scalar or (fx, fy);
    c None (image centre) or (cx, cy) -- the arguments of util.gen_rays (src/util/util.py:238-276)."""
    poses = _f32c(poses)
    if poses.dim() != 3 or poses.shape[1] not in (3, 4) or poses.shape[2] != 4:
        raise ValueError(f"poses must be (B, 3|4, 4), got {tuple(poses.shape)}")
    f = torch.as_tensor(focal, dtype=torch.float32).reshape(-1)
    fx, fy = float(f[0]), float(f[-1])
    if c is None:
        cx, cy = width * 0.5, height * 0.5
    else:
        c = torch.as_tensor(c).reshape(-1)
        cx, cy = float(c[0]), float(c[1])
    return Frames(_ptr(poses), poses.shape[1] * 4, poses.shape[0], int(width), int(height), fx, fy, cx, cy,
                  float(z_near), float(z_far)), poses


def render_rays(cam_bufs, hl, wl, mlp_coarse, mlp_fine, rays, n_coarse, n_imp, n_depth, depth_std, lindisp,
                white_bkgd, seed=0, draws=None, frames=None, first_ray=0, n_rays=None):
    """The whole no-grad render in one C call (pn_render_rays): mlp_* = (packed, tables) pairs of the tensor-core
    chain, mlp_fine None = the fine pass uses mlp_coarse.  rays (R,8) ordered (object, ray).  Returns
    (weights, rgb, depth, z) for the coarse pass and the same tuple (or None) for the fine pass.
    With ``frames`` (a frames_struct pair) instead of rays: pn_render_frames on pixels [first_ray, first_ray + n_rays)
    of the frames -- the rays are generated inside the kernels."""
    if frames is None:
        rays = _f32c(rays)
        r = rays.shape[0]
        dev = rays.device
    else:
        r, dev = int(n_rays), frames[1].device
    cam = Camera(_ptr(cam_bufs["poses"]), _ptr(cam_bufs["focal"]), _ptr(cam_bufs["c"]),
                 cam_bufs["scale_x"], cam_bufs["scale_y"], hl, wl, cam_bufs["n_objects"], cam_bufs["n_views"])
    cfg = RenderConfig(n_coarse, n_imp, n_depth, float(depth_std), int(bool(lindisp)), int(bool(white_bkgd)), seed)
    d = {k: _f32c(v) for k, v in (draws or {}).items()}
    dr = RenderDraws(_ptr(d.get("u_coarse")), _ptr(d.get("u_bin")), _ptr(d.get("u_in")), _ptr(d.get("noise")))
    n_tex = mlp_coarse[1].shape[1]
    for pk, tb in (mlp_coarse,) + ((mlp_fine,) if mlp_fine is not None else ()):
        if tb.dtype != torch.float16 or tb.dim() != 3 or tb.shape[1] != n_tex or not tb.is_contiguous():
            raise RuntimeError("render_rays needs contiguous fp16 lin_z tables of shape (3, texels, 512)")
    mc = RenderMlp(mlp_coarse[0].data_ptr(), mlp_coarse[1].data_ptr())
    mf = RenderMlp(mlp_fine[0].data_ptr(), mlp_fine[1].data_ptr()) if mlp_fine is not None else None

    def outs(k):
        t = (torch.empty(r, k, device=dev), torch.empty(r, 3, device=dev), torch.empty(r, device=dev),
             torch.empty(r, k, device=dev))
        return t, RenderOut(_ptr(t[1]), _ptr(t[2]), _ptr(t[0]), _ptr(t[3]))

    k_f = n_imp + n_depth
    coarse, oc = outs(n_coarse)
    fine, of = outs(n_coarse + k_f) if k_f > 0 else (None, None)
    nbytes = lib().pn_render_rays_workspace_bytes(r, cam.n_views, C.byref(cfg))
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    ev, keep = None, None
    if PROFILE is not None:
        ev, keep = RenderEvents(), {}
        for name, key in (("taps", "features"), ("mlp", "mlp"), ("composite", "composite"), ("sample", "sampling")):
            for p in range(2 if k_f > 0 else 1):
                pair = (_new_event(), _new_event())
                getattr(ev, name)[p][0], getattr(ev, name)[p][1] = pair[0].cuda_event, pair[1].cuda_event
                PROFILE.setdefault(key, []).append(pair)
    if frames is not None:
        _check(lib().pn_render_frames(C.byref(cam), C.byref(mc), C.byref(mf) if mf is not None else None, n_tex,
                                      C.byref(frames[0]), int(first_ray), r, _ptr(_linspace(n_coarse, dev)), C.byref(cfg),
                                      C.byref(dr), C.byref(oc), C.byref(of) if of is not None else None, ws.data_ptr(),
                                      nbytes, C.byref(ev) if ev is not None else None, _stream()), "pn_render_frames")
        return coarse, fine
    _check(lib().pn_render_rays(C.byref(cam), C.byref(mc), C.byref(mf) if mf is not None else None, n_tex, _ptr(rays),
                                r, _ptr(_linspace(n_coarse, dev)), C.byref(cfg), C.byref(dr), C.byref(oc),
                                C.byref(of) if of is not None else None, ws.data_ptr(), nbytes,
                                C.byref(ev) if ev is not None else None, _stream()), "pn_render_rays")
    return coarse, fine


def gen_rays(poses, width, height, focal, z_near, z_far, c=None):
    """util.gen_rays (src/util/util.py:238-276) on the device: poses (B, 4, 4) or (B, 3, 4) -> (B, H, W, 8)."""
    poses = _f32c(poses)
    if poses.dim() != 3 or poses.shape[1] not in (3, 4) or poses.shape[2] != 4:
        raise ValueError(f"poses must be (B, 3|4, 4), got {tuple(poses.shape)}")
    f = torch.as_tensor(focal, dtype=torch.float32).reshape(-1)
    fx, fy = float(f[0]), float(f[-1])
    if c is None:
        cx, cy = width * 0.5, height * 0.5
    else:
        c = torch.as_tensor(c).reshape(-1)
        cx, cy = float(c[0]), float(c[1])
    n = poses.shape[0]
    rays = torch.empty(n, height, width, 8, device=poses.device, dtype=torch.float32)
    with _Timed("gen_rays"):
        _check(lib().pn_gen_rays(_ptr(poses), poses.shape[1] * 4, n, width, height, fx, fy, cx, cy,
                                 float(z_near), float(z_far), _ptr(rays), _stream()), "pn_gen_rays")
    return rays


def encoder_latent(levels, dtype=torch.float32):
    """SpatialEncoder.forward's tail (encoder.py:150-163) in one kernel: list of NCHW fp32 level maps ->
    channels-last (V, H0, W0, sum C) latent = cat([interpolate(l, (H0, W0), bilinear, align_corners=True)])."""
    levels = [_f32c(l) for l in levels]
    n = len(levels)
    v = levels[0].shape[0]
    dev = levels[0].device
    if any(l.dim() != 4 or l.shape[0] != v or not l.is_cuda for l in levels):
        raise RuntimeError("encoder_latent needs CUDA level maps of shape (V, C_l, H_l, W_l)")
    ctot = sum(l.shape[1] for l in levels)
    out = torch.empty(v, levels[0].shape[2], levels[0].shape[3], ctot, device=dev, dtype=dtype)
    maps = (_vp * n)(*[l.data_ptr() for l in levels])
    ch = (_int * n)(*[l.shape[1] for l in levels])
    hh = (_int * n)(*[l.shape[2] for l in levels])
    ww = (_int * n)(*[l.shape[3] for l in levels])
    code = DTYPE_F32 if dtype == torch.float32 else DTYPE_F16
    with _Timed("encoder_latent"):
        _check(lib().pn_encoder_latent(maps, ch, hh, ww, n, v, out.data_ptr(), code, _stream()), "pn_encoder_latent")
    return out


def frame_to_u8(rgb):
    """(frames * 255).astype(np.uint8) of eval/gen_video.py:234-236, on the device; same shape, uint8."""
    rgb = _f32c(rgb)
    out = torch.empty(rgb.shape, device=rgb.device, dtype=torch.uint8)
    with _Timed("frame_to_u8"):
        _check(lib().pn_frame_to_u8(_ptr(rgb), _ptr(out, torch.uint8), rgb.numel(), _stream()), "pn_frame_to_u8")
    return out


_FMT = {torch.float16: 0, torch.bfloat16: 1}
_ERR_WORD = {}


def _err_word(dev):
    key = str(dev)
    if key not in _ERR_WORD:
        _ERR_WORD[key] = torch.zeros(64, device=dev, dtype=torch.int32)
    return _ERR_WORD[key]


def gemm_tc_rows(a, b, mask=None, acc32=None, accumulate=False, want16=True, scale=1.0, resid=None, colsum=None,
                 colsum2=None):
    """scale * a[rows, k] @ b[k, n] on the tensor cores (16-bit operands, fp32 accumulate), optionally masked by
    the relu output `mask`, added to `acc32` (fp32, in place) and/or `resid` (bf16); `colsum` / `colsum2` (n,) fp32
    receive += the column sums of the result.  Returns (acc32 or None, bf16 copy or None)."""
    rows, k = a.shape
    n = b.shape[1]
    assert b.shape[0] == k and a.is_contiguous() and b.is_contiguous()
    out16 = torch.empty(rows, n, device=a.device, dtype=torch.bfloat16) if want16 else None
    with _Timed("gemm_rows"):
        _check(lib().pn_gemm_tc_rows(_ptr(a, a.dtype), _FMT[a.dtype], _ptr(b, b.dtype), _FMT[b.dtype], rows, k, n,
                                     _ptr(mask, mask.dtype if mask is not None else None), _ptr(acc32),
                                     int(bool(accumulate)), _ptr(resid, torch.bfloat16), _ptr(out16, torch.bfloat16),
                                     float(scale), _ptr(colsum), _ptr(colsum2), _err_word(a.device).data_ptr(),
                                     _stream()), "pn_gemm_tc_rows")
    return acc32, out16


def gemm_tc_rows_slabs(a_slabs, b, n_rows, acc32=None, accumulate=False, scale=1.0):
    """sum_s a_slabs[s, :n_rows] @ b[s*512:(s+1)*512] in one launch: a_slabs (S, slab_rows, 512) 16-bit with
    slab_rows % 256 == 0, b (S*512, n) of the same format; fp32 result (n_rows, n)."""
    s_, slab_rows, _ = a_slabs.shape
    n = b.shape[1]
    assert b.shape[0] == s_ * 512 and a_slabs.is_contiguous() and b.is_contiguous() and a_slabs.dtype == b.dtype
    if acc32 is None:
        acc32 = torch.empty(n_rows, n, device=a_slabs.device, dtype=torch.float32)
    with _Timed("gemm_rows"):
        _check(lib().pn_gemm_tc_rows_slabs(_ptr(a_slabs, a_slabs.dtype), slab_rows, _FMT[a_slabs.dtype],
                                           _ptr(b, b.dtype), n_rows, s_ * 512, n, _ptr(acc32), int(bool(accumulate)),
                                           float(scale), _err_word(a_slabs.device).data_ptr(), _stream()),
               "pn_gemm_tc_rows_slabs")
    return acc32


def gemm_tc_wgrad(g, x, out, n_store=None):
    """out[512, :n_store] += g[rows, 512]^T @ x[rows, 512 | 64] on the tensor cores."""
    rows = g.shape[0]
    n_store = x.shape[1] if n_store is None else n_store
    assert out.dim() == 2 and out.shape[0] == 512 and out.is_contiguous()
    with _Timed("gemm_wgrad"):
        _check(lib().pn_gemm_tc_wgrad(_ptr(g, g.dtype), _FMT[g.dtype], g.shape[1], _ptr(x, x.dtype), _FMT[x.dtype],
                                      x.shape[1], rows, n_store, _ptr(out), out.shape[1],
                                      _err_word(g.device).data_ptr(), _stream()), "pn_gemm_tc_wgrad")
    return out


def mlp_forward_tc_train(packed, tables, feat_x, n_objects, n_views, ppo):
    """Tensor-core forward that keeps the bf16 operand of every 512-wide GEMM: returns (out (P,4), acts)."""
    n_rows = n_objects * n_views * ppo
    dev = feat_x.device
    if feat_x.shape != (n_rows, 64) or feat_x.dtype != torch.float16 or not feat_x.is_contiguous():
        raise RuntimeError("tensor-core MLP needs contiguous fp16 feat_x of shape (rows, 64)")
    out = torch.empty(n_objects * ppo, 4, device=dev, dtype=torch.float32)
    nbytes = lib().pn_mlp_tc_workspace_bytes(n_rows, n_views)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    abytes = lib().pn_mlp_tc_acts_bytes(n_rows)
    acts = torch.empty(abytes, device=dev, dtype=torch.uint8)
    with _Timed("mlp_train_fwd"):
        _check(lib().pn_mlp_forward_tc_train(packed.data_ptr(), tables.data_ptr(), tables.shape[1],
                                             _ptr(feat_x, torch.float16), n_objects, n_views, ppo, _ptr(out), None,
                                             ws.data_ptr(), nbytes, acts.data_ptr(), abytes, _stream()),
               "pn_mlp_forward_tc_train")
    return out, acts


def tc_acts_view(acts, n_rows):
    """The saved-operand buffer of mlp_forward_tc_train as a (11, n_rows, 512) bf16 view (slot pitch = n_rows
    rounded up to whole 128-row tiles)."""
    pitch = (n_rows + 127) // 128 * 128
    return acts[: 11 * pitch * 1024].view(torch.bfloat16).reshape(11, pitch, 512)[:, :n_rows]


def pack_weights_bf16(w):
    """bf16 copies of the thirteen 512x512 matrices for the tensor-core backward."""
    s, keep = pack_weights_fp32(w)
    nbytes = lib().pn_mlp_tc_w16_bytes()
    w16 = torch.empty(nbytes, device=keep["lin_in.weight"].device, dtype=torch.uint8)
    _check(lib().pn_prepare_weights_bf16(C.byref(s), w16.data_ptr(), nbytes, _stream()), "pn_prepare_weights_bf16")
    return w16


def mlp_backward_tc(packed32, w16, feat_x, latent_nhwc, n_objects, n_views, ppo, acts, out, g_out, grads,
                    want_g_feat_z, want_g_h0=False):
    """grads: dict keyed by weight_names(), accumulated into (+=); returns dL/d feat_z (rows, 512) fp32 or None;
    with want_g_h0 also dL/dh_0 (rows, 512) bf16, a view into the call's workspace."""
    s, _keep = packed32
    sg = _fill_struct(MlpGrads(), grads)
    n_rows = n_objects * n_views * ppo
    dev = feat_x.device
    g_fz = torch.empty(n_rows, D_LATENT, device=dev, dtype=torch.float32) if want_g_feat_z else None
    nbytes = lib().pn_mlp_backward_tc_workspace_bytes(n_rows, n_views)
    ws = torch.empty(nbytes, device=dev, dtype=torch.uint8)
    with _Timed("mlp_train_bwd"):
        _check(lib().pn_mlp_backward_tc(C.byref(s), w16.data_ptr(), _ptr(feat_x, torch.float16), _ptr(latent_nhwc),
                                        n_objects, n_views, ppo, acts.data_ptr(), _ptr(out), _ptr(_f32c(g_out)),
                                        C.byref(sg), _ptr(g_fz), ws.data_ptr(), nbytes, _stream()),
               "pn_mlp_backward_tc")
    if want_g_h0:
        off = lib().pn_mlp_backward_tc_gh0_offset(n_rows, n_views)
        return g_fz, ws[off:off + n_rows * D_HIDDEN * 2].view(torch.bfloat16).view(n_rows, D_HIDDEN)
    return g_fz


# every entry point that launches kernels runs with its tensors' device current (see _on_device)
for _name in ("sample_coarse", "sample_fine", "prepare_latent", "point_features", "mlp_forward_fp32",
              "mlp_forward_fp32_train", "mlp_backward_fp32", "point_features_bwd", "depth_sample_bwd", "point_taps",
              "prepare_latent_tables", "pack_weights_tc", "mlp_forward_tc", "composite_fwd", "composite_bwd",
              "point_taps_grouped", "mlp_forward_tc_grouped", "prepare_latent_tables_f32", "pack_weights_x3",
              "point_taps_f32", "mlp_forward_tc_x3", "render_rays", "gen_rays", "encoder_latent", "frame_to_u8", "gemm_tc_rows", "gemm_tc_rows_slabs",
              "gemm_tc_wgrad", "mlp_forward_tc_train", "pack_weights_bf16", "mlp_backward_tc"):
    globals()[_name] = _on_device(globals()[_name])
del _name
