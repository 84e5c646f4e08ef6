"""TalkingHeadsBlock (models/layers/attentions/talking_heads.py:5-14): learned H x H mixing across
heads, `einsum('h i, b h ... -> b i ...')`, orthogonal init, parameter `talking_heads_transform`.

The reference applies it to the materialised [B,H,Nq,Nk] logits / probabilities
(attention.py:44-52).  Here the module owns the parameter (same name, same auto-numbering under the
attention block: TalkingHeadsBlock_0 = pre-softmax, TalkingHeadsBlock_1 = post-softmax) and hands
it to the row kernel that does mix -> softmax -> mix in one pass over the logits (attn_mat.cu).
"""
from sav import linen as nn


class TalkingHeadsBlock(nn.Module):
    num_heads: int

    @nn.compact
    def __call__(self, inputs=None):
        transform = self.param('talking_heads_transform', nn.initializers.orthogonal(),
                               (self.num_heads, self.num_heads))
        if inputs is not None:
            raise NotImplementedError("TalkingHeadsBlock on a free-standing tensor: it is fused into "
                                      "AttentionBlock(talking_heads=True)")
        return transform
