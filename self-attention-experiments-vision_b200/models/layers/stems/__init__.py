from models.layers.stems.patch_embed import PatchEmbedBlock  # noqa: F401
