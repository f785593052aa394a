from .resnet import ResNet, Bottleneck, resnet50

__all__ = ["ResNet", "Bottleneck", "resnet50"]
