from .pyramidnet import PyramidNet, PyramidBlock, pyramidnet, pyramid_widths
from .mnist import MnistCNN, MLP

__all__ = ["PyramidNet", "PyramidBlock", "pyramidnet", "pyramid_widths", "MnistCNN", "MLP"]
