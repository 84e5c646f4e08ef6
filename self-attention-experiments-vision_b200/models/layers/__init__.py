from models.layers.attentions import AttentionBlock, SelfAttentionBlock, TalkingHeadsBlock  # noqa: F401
from models.layers.attentions import CvTAttentionBlock, CvTSelfAttentionBlock  # noqa: F401
from models.layers.feedforwards import FFBlock, LeFFBlock  # noqa: F401
from models.layers.normalizations import LayerScaleBlock  # noqa: F401
from models.layers.position_embed import (AddAbsPosEmbed, FixedPositionalEmbedding,  # noqa: F401
                                           RotaryPositionalEmbedding)
from models.layers.regularization import StochasticDepthBlock  # noqa: F401
from models.layers.stems import Image2TokenBlock, PatchEmbedBlock  # noqa: F401
from models.layers.squeeze_excite import SqueezeExciteBlock  # noqa: F401
