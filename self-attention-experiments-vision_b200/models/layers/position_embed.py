"""AddAbsPosEmbed (models/layers/position_embed.py:48-57): learned [1,N,C] table, N(0,0.02) init.

`cls` may be passed so that the class-token concat (vit.py:81-85) and the position add are one
elementwise kernel.  The reference's rotary helpers (position_embed.py:8-45) are unfinished and
unused by every model; they are not provided (out of scope).
"""
from sav import functional as F
from sav import linen as nn


class AddAbsPosEmbed(nn.Module):
    embed_init: object = None

    @nn.compact
    def __call__(self, inputs, cls=None):
        assert inputs.ndim == 3
        n = inputs.shape[1] + (1 if cls is not None else 0)
        pos = self.param('pos_embed', self.embed_init or nn.initializers.normal(stddev=0.02),
                         (1, n, inputs.shape[2]))
        return F.embed(inputs, cls, pos)
