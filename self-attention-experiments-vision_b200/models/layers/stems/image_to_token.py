"""Image2TokenBlock (models/layers/stems/image_to_token.py:8-48), CeiT's convolutional stem:
Conv 7x7 / stride 2 with explicit padding = the patch shape (im2col + tcgen05 GEMM on the raw images)
-> BatchNorm -> max_pool 3x3 / stride 2 ('VALID', flax's default) -> 4x4 patches of the feature map
-> Dense.  Names: Conv_0/kernel, BatchNorm_0/{scale,bias} (+ batch_stats), Dense_0/kernel."""
from typing import Tuple

from sav import functional as F
from sav import linen as nn


class Image2TokenBlock(nn.Module):
    patch_shape: Tuple[int, int] = (4, 4)
    num_ch: int = 32
    conv_kernel_size: int = 7
    conv_stride: int = 2
    pool_window_size: int = 3
    pool_stride: int = 2
    embed_dim: int = 192
    use_bias: bool = False
    bn_momentum: float = 0.9
    bn_epsilon: float = 1e-5
    dtype: object = None

    @nn.compact
    def __call__(self, inputs, is_training: bool):
        if self.pool_window_size != 3 or self.pool_stride != 2:
            raise NotImplementedError("Image2TokenBlock: max_pool 3x3 / stride 2 only (every CeiT config)")
        k, s = self.conv_kernel_size, self.conv_stride
        x = nn.Conv(features=self.num_ch, use_bias=self.use_bias, kernel_size=(k, k), strides=(s, s),
                    padding=[(self.patch_shape[0],) * 2, (self.patch_shape[1],) * 2])(inputs)   # :24-29
        x = nn.BatchNorm(use_running_average=not is_training, momentum=self.bn_momentum,
                         epsilon=self.bn_epsilon)(x)                                           # :30-33
        x = F.maxpool3x3s2(x, valid=True)                                                      # :34-38
        x = F.patchify_act(x, self.patch_shape[0], self.patch_shape[1])                        # :39-44
        return nn.Dense(features=self.embed_dim, use_bias=self.use_bias)(x)                    # :46-48
