from .pyramidnet import PyramidNet, PyramidBlock, pyramidnet, pyramid_widths
from .mnist import MnistCNN, MLP
from .resnet import Bottleneck, ResNet50, StemConv7x7s2, resnet50

__all__ = ["PyramidNet", "PyramidBlock", "pyramidnet", "pyramid_widths", "MnistCNN", "MLP", "ResNet50", "resnet50", "Bottleneck", "StemConv7x7s2"]
