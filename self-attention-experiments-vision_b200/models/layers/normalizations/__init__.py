from models.layers.normalizations.layerscale import LayerScaleBlock  # noqa: F401
