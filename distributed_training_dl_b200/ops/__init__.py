from . import functional
from .modules import (AvgPool2d, BatchNorm2d, Conv2d, Flatten, GlobalAvgPool2d, InputLayout, Linear,
                      MaxPool2d, OutputLayout, pack_weights, packed_layers, set_precision)

__all__ = ["functional", "AvgPool2d", "BatchNorm2d", "Conv2d", "Flatten", "GlobalAvgPool2d",
           "InputLayout", "Linear", "MaxPool2d", "OutputLayout", "pack_weights", "packed_layers", "set_precision"]
