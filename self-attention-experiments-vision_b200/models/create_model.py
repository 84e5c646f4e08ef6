"""create_model(model_name, num_classes=1000, dtype=float32) -- models/create_model.py:6-215.

Names and hyper-parameters are the reference's; `deit_ti_patch16` / `deit_s_patch16` are added
(the reference has no DeiT constructor; DeiT-Ti/S are ViT(12 layers, 3/6 heads, 192/384)).
`dtype` is accepted for signature compatibility; the compute path is always bf16 storage with fp32
accumulation / statistics.
"""
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


def create_model(model_name: str, num_classes: int = 1000, dtype=None):
    if model_name in _VIT:
        layers, heads, dim, p = _VIT[model_name]
        return ViT(num_classes=num_classes, num_layers=layers, num_heads=heads, embed_dim=dim,
                   patch_shape=(p, p), dtype=dtype)
    raise NotImplementedError(f'{model_name} is not built yet (see DESIGN.md status table)')
