from models.vit import ViT  # noqa: F401
from models.cait import CaiT  # noqa: F401
from models.cvt import ConvTokenEmbedBlock, CvT, Stage, StageBlock  # noqa: F401
from models.botnet import BoTBlock, BoTMHSA, BoTNet, BottleneckResNetBlock, RelativeLogits  # noqa: F401
from models.mlp_mixer import MLPMixer, MixerBlock  # noqa: F401
from models.tnt import TNT  # noqa: F401
from models.ceit import CeiT  # noqa: F401
from models.create_model import create_model  # noqa: F401
