from models.layers.feedforwards.ff import FFBlock  # noqa: F401
from models.layers.feedforwards.leff import LeFFBlock  # noqa: F401
