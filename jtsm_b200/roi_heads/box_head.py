"""`DiscriminativeAdaptionNeck`: the WSL box head (fc6 / fc7 of VGG16: flatten -> [Linear, ReLU, Dropout(0.5)] per fc dim)
with the reference's layer names, initialisation and `output_shape`
(projects/WSL/wsl/modeling/roi_heads/box_head.py:18-107).  The GEMMs stay in torch (north_star); the class exists so
that a checkpoint of the reference loads into `OICRROIHeadsB200` key for key (`box_head.fc1.weight`, ...)."""
from types import SimpleNamespace

import numpy as np
import torch
from torch import nn


class DiscriminativeAdaptionNeck(nn.Sequential):
    def __init__(self, input_shape, *, conv_dims=(), fc_dims=(4096, 4096), conv_norm=""):
        super().__init__()
        if len(conv_dims):
            raise NotImplementedError("conv layers in the box head are not used by the WSL configs (NUM_CONV: 0)")
        assert len(fc_dims) > 0
        self._output_size = (input_shape.channels, input_shape.height, input_shape.width)
        self.fcs = []
        for k, fc_dim in enumerate(fc_dims):
            if k == 0:
                self.add_module("flatten", nn.Flatten())
            fc = nn.Linear(int(np.prod(self._output_size)), fc_dim)
            self.add_module("fc{}".format(k + 1), fc)
            self.add_module("fc_relu{}".format(k + 1), nn.ReLU(inplace=True))
            self.add_module("fc_dropout{}".format(k + 1), nn.Dropout(p=0.5, inplace=False))
            self.fcs.append(fc)
            self._output_size = fc_dim
        for layer in self.fcs:
            torch.nn.init.normal_(layer.weight, std=0.005)
            torch.nn.init.constant_(layer.bias, 0.1)

    def forward(self, x):
        for layer in self:
            x = layer(x)
        return x

    @property
    def output_shape(self):
        o = self._output_size
        if isinstance(o, int):
            return SimpleNamespace(channels=o, height=None, width=None, stride=None)
        return SimpleNamespace(channels=o[0], height=o[1], width=o[2], stride=None)
