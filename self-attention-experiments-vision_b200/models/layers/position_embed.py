"""Position embeddings (models/layers/position_embed.py).

AddAbsPosEmbed (position_embed.py:48-57): learned [1,N,C] table, N(0,0.02) init; `cls` may be
passed so that the class-token concat (vit.py:81-85) and the position add are one elementwise
kernel.

Rotary helpers (position_embed.py:8-45): present in the reference's export list
(models/layers/__init__.py:4) but unused by every model and unfinished -- `apply_rotary_pos_emb`
unpacks (sin, cos) from what `FixedPositionalEmbedding` returns as ONE frequency tensor, and
`FixedPositionalEmbedding` reads a `dtype` field it does not declare.  Built here with
the intended RoFormer reading -- sin / cos of the frequency table, pairs (2i, 2i+1) rotated -- and the
reference's literal float32 table inv_freq = 1 / (10e4**arange(0, dim, 2) / dim).  Parity unpinned.
"""
import torch

from sav import functional as F
from sav import linen as nn
from sav.tape import Var


def rotate_every_two(x):
    """position_embed.py:8-14 on a plain tensor: (x0, x1, x2, x3, ...) -> (-x1, x0, -x3, x2, ...).
    Host-side helper for tests / table construction; the device path is F.rotary."""
    x1 = x[..., ::2]
    x2 = x[..., 1::2]
    return torch.stack((-x2, x1), dim=-1).reshape(x.shape)


def apply_rotary_pos_emb(x, sincos=None, seq_dim=0):
    """position_embed.py:17-20: x cos + rotate_every_two(x) sin.  `sincos` is accepted for signature
    compatibility; the sin / cos tables are evaluated inside the kernel from the position index."""
    return F.rotary(x, seq_dim)


class FixedPositionalEmbedding(nn.Module):
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, seq_dim=0):
        from sav import ops
        dim = inputs.shape[-1]
        n = inputs.shape[seq_dim]
        dev = inputs.v.device
        if dev.type == "meta":
            return Var(torch.empty((n, dim), dtype=torch.float32, device="meta"), requires_grad=False)
        return Var(ops.rotary_table(n, dim, dev), requires_grad=False)


class RotaryPositionalEmbedding(FixedPositionalEmbedding):
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, seq_dim=0):
        return apply_rotary_pos_emb(inputs, None, seq_dim)


class AddAbsPosEmbed(nn.Module):
    embed_init: object = None

    @nn.compact
    def __call__(self, inputs, cls=None):
        assert inputs.ndim == 3
        n = inputs.shape[1] + (1 if cls is not None else 0)
        pos = self.param('pos_embed', self.embed_init or nn.initializers.normal(stddev=0.02),
                         (1, n, inputs.shape[2]))
        return F.embed(inputs, cls, pos)
