"""PatchEmbedBlock (models/layers/stems/patch_embed.py:8-26).

'b (h ph) (w pw) c -> b (h w) (ph pw c)' is a coalesced patchify kernel that also casts to bf16
(and, given `sav.HWCN(images)`, absorbs train.py:80's 'H W C N -> N H W C'); the Dense is a GEMM.
"""
from typing import Tuple

from sav import functional as F
from sav import linen as nn
from sav.train import HWCN


class PatchEmbedBlock(nn.Module):
    patch_shape: Tuple[int] = (16, 16)
    embed_dim: int = 0
    use_bias: bool = False
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, *unused_args, **unused_kwargs):
        ph, pw = self.patch_shape
        if isinstance(inputs, HWCN):
            x = F.patchify(inputs.images, ph, pw, hwcn=True)
        else:
            x = F.patchify(inputs, ph, pw, hwcn=False)
        return nn.Dense(features=self.embed_dim, use_bias=self.use_bias)(x)
