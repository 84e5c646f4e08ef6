from models.layers.regularization.stochastic_depth import StochasticDepthBlock  # noqa: F401
