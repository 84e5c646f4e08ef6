from models.layers.attentions.talking_heads import TalkingHeadsBlock  # noqa: F401
from models.layers.attentions.attention import AttentionBlock, SelfAttentionBlock  # noqa: F401
from models.layers.attentions.cvt_attention import (ConvProjectionBlock, CvTAttentionBlock,  # noqa: F401
                                                    CvTSelfAttentionBlock)
