from models.layers.attentions.attention import AttentionBlock, SelfAttentionBlock  # noqa: F401
