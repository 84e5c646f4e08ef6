from models.layers.stems.patch_embed import PatchEmbedBlock  # noqa: F401
from models.layers.stems.image_to_token import Image2TokenBlock  # noqa: F401
