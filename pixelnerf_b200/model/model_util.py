"""String-typed factories, reference src/model/model_util.py."""
from .encoder import SpatialEncoder
from .resnetfc import ResnetFC


def make_mlp(conf, d_in, d_latent=0, allow_empty=False, **kwargs):
    mlp_type = conf.get_string("type", "mlp")
    if mlp_type == "resnet":
        return ResnetFC.from_conf(conf, d_in, d_latent=d_latent, **kwargs)
    if mlp_type == "empty" and allow_empty:
        return None
    if mlp_type == "mlp":
        raise NotImplementedError("type=mlp (ImplicitNet) is outside the B200 hot path; no shipped conf selects it")
    raise NotImplementedError("Unsupported MLP type")


def make_encoder(conf, **kwargs):
    enc_type = conf.get_string("type", "spatial")
    if enc_type == "spatial":
        return SpatialEncoder.from_conf(conf, **kwargs)
    if enc_type == "global":
        raise NotImplementedError("global ImageEncoder is outside the B200 hot path (use_global_encoder=False everywhere)")
    raise NotImplementedError("Unsupported encoder type")
