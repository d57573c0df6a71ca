"""Sub-pixel convolution super-resolution CNN.

Parity target: ``Net(upscale_factor=2)`` from the absent ``models.sr_4k_2x``
(Fairscale-DDP.py:13,74): an image-to-image CNN that upsamples by ``upscale_factor``.
The source is not in the reference; this is the standard efficient sub-pixel design
(conv stack → ``PixelShuffle``) with RGB in / RGB out, which satisfies the script's
contract ``model(lr_batch) -> hr_batch`` trained with ``nn.MSELoss``.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..ops.layers import Conv2d

__all__ = ["Net"]


class Net(nn.Module):
    def __init__(self, upscale_factor: int = 2, channels: int = 3, width: int = 64):
        super().__init__()
        self.relu = nn.ReLU(inplace=True)
        self.conv1 = Conv2d(channels, width, 5, 1, 2)
        self.conv2 = Conv2d(width, width, 3, 1, 1)
        self.conv3 = Conv2d(width, width // 2, 3, 1, 1)
        self.conv4 = Conv2d(width // 2, channels * upscale_factor ** 2, 3, 1, 1)
        self.pixel_shuffle = nn.PixelShuffle(upscale_factor)
        for m, gain in ((self.conv1, "relu"), (self.conv2, "relu"), (self.conv3, "relu"), (self.conv4, "linear")):
            nn.init.orthogonal_(m.weight, nn.init.calculate_gain(gain))
            nn.init.zeros_(m.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.relu(self.conv1(x))
        x = self.relu(self.conv2(x))
        x = self.relu(self.conv3(x))
        return self.pixel_shuffle(self.conv4(x))
