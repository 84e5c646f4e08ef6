"""create_model(model_name, num_classes=1000, dtype=float32) -- models/create_model.py:6-215.

Names and hyper-parameters are the reference's; `deit_ti_patch16` / `deit_s_patch16` are added
(the reference has no DeiT constructor; DeiT-Ti/S are ViT(12 layers, 3/6 heads, 192/384)).
`dtype` is accepted for signature compatibility; the compute path is always bf16 storage with fp32
accumulation / statistics.
"""
from models.botnet import BoTNet
from models.cait import CaiT
from models.cvt import CvT
from models.mlp_mixer import MLPMixer
from models.tnt import TNT
from models.ceit import CeiT
from models.vit import ViT

_VIT = {
    # name: (num_layers, num_heads, embed_dim, patch)           create_model.py:10-37
    'vit_b_patch32': (12, 12, 768, 32),
    'vit_b_patch16': (12, 12, 768, 16),
    'vit_l_patch32': (24, 16, 1024, 32),
    'vit_l_patch16': (24, 16, 1024, 16),
    'deit_ti_patch16': (12, 3, 192, 16),
    'deit_s_patch16': (12, 6, 384, 16),
}


_CAIT = {
    # name: (num_layers, num_heads, embed_dim, stoch_depth_rate, layerscale_eps)  create_model.py:79-168
    'cait_xxs_24': (24, 4, 192, 0.05, 1e-5),
    'cait_xxs_36': (36, 4, 192, 0.1, 1e-6),
    'cait_xs_24': (24, 6, 288, 0.05, 1e-5),
    'cait_xs_36': (36, 6, 288, 0.1, 1e-6),
    'cait_s_24': (24, 8, 384, 0.1, 1e-6),
    'cait_s_36': (36, 8, 384, 0.2, 1e-6),
    'cait_s_48': (48, 8, 384, 0.3, 1e-6),
    'cait_m_24': (24, 16, 768, 0.2, 1e-5),
    'cait_m_36': (36, 16, 768, 0.3, 1e-6),
    'cait_m_48': (48, 16, 768, 0.4, 1e-6),
}


_BOTNET = {
    # name: stage_sizes                                          create_model.py:38-49
    'botnet_t3': (3, 4, 6, 6),
    'botnet_t4': (3, 4, 23, 6),
    'botnet_t5': (3, 4, 23, 12),
}

_CVT = {
    # name: (stage_sizes, num_heads, embed_dim)                  create_model.py:169-183
    # the reference writes embed_dim (64, 192, 368) for cvt-13 / cvt-21; 368 % 6 heads != 0 trips its
    # own assert (cvt_attention.py:65), cvt_test.py:14-15 and the CvT paper use 384
    'cvt-13': ((1, 2, 10), (1, 3, 6), (64, 192, 384)),
    'cvt-21': ((1, 4, 16), (1, 3, 6), (64, 192, 384)),
    'cvt-w24': ((2, 2, 20), (3, 12, 16), (192, 768, 1024)),
}


# create_model.py:50-63: (inner heads, outer heads, inner dim, outer dim), 12 layers
_TNT = {'tnt_s_patch16': (4, 10, 40, 640), 'tnt_b_patch16': (4, 6, 24, 384)}

# create_model.py:64-78: (heads, embed_dim), 12 layers
_CEIT = {'ceit_t': (3, 192), 'ceit_s': (6, 384), 'ceit_b': (12, 768)}

_MIXER = {
    # create_model.py:184-213 verbatim (layers, embed_dim, patch).  The reference spells the B/16
    # entry 'mixer_s_patch32' a second time (create_model.py:199), so it is unreachable there; it is
    # offered here under the name the pattern implies.  mixer_l_patch16 has 32 layers in
    # create_model.py:209-213 against 24 in mlp_mixer_test.py:19: the named surface wins.
    'mixer_s_patch32': (8, 512, 32), 'mixer_s_patch16': (8, 512, 16),
    'mixer_b_patch32': (12, 768, 32), 'mixer_b_patch16': (12, 768, 16),
    'mixer_l_patch32': (24, 1024, 32), 'mixer_l_patch16': (32, 1024, 16),
}


def create_model(model_name: str, num_classes: int = 1000, dtype=None):
    if model_name in _VIT:
        layers, heads, dim, p = _VIT[model_name]
        return ViT(num_classes=num_classes, num_layers=layers, num_heads=heads, embed_dim=dim,
                   patch_shape=(p, p), dtype=dtype)
    if model_name in _CAIT:
        layers, heads, dim, sd, eps = _CAIT[model_name]
        return CaiT(num_classes=num_classes, num_layers=layers, num_layers_token_only=2,
                    num_heads=heads, embed_dim=dim, patch_shape=(16, 16), stoch_depth_rate=sd,
                    layerscale_eps=eps, dtype=dtype)
    if model_name in _BOTNET:
        return BoTNet(num_classes=num_classes, stage_sizes=_BOTNET[model_name], dtype=dtype)
    if model_name in _CVT:
        sizes, heads, dims = _CVT[model_name]
        return CvT(num_classes=num_classes, stage_sizes=sizes, num_heads=heads, embed_dim=dims)
    if model_name in _MIXER:
        layers, dim, p = _MIXER[model_name]
        return MLPMixer(num_classes=num_classes, num_layers=layers, embed_dim=dim, patch_shape=(p, p))
    if model_name in _CEIT:
        heads, dim = _CEIT[model_name]
        return CeiT(num_classes=num_classes, num_layers=12, num_heads=heads, embed_dim=dim)
    if model_name in _TNT:
        ih, oh, idim, odim = _TNT[model_name]
        return TNT(num_classes=num_classes, num_layers=12, inner_num_heads=ih, outer_num_heads=oh,
                   inner_embed_dim=idim, outer_embed_dim=odim)
    raise RuntimeError('Model not found.')                  # create_model.py:214-215
