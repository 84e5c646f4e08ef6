from models.layers.feedforwards.ff import FFBlock  # noqa: F401
