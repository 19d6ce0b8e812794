"""The fedb200 model zoo: the reference's 18 model files / 44 constructor configurations
 rebuilt on ``fedb200.ops.layers`` with identical ``state_dict`` layouts.

``REGISTRY`` maps a name (as a user would type after ``--model``) to a zero-argument factory;
``build(name)`` instantiates it.  The reference selects its model by editing source
(src/main.py:63-77, src/server.py:158); here it is a flag.
"""
from .lenet import LeNet
from .mobilenet import MobileNet
from .mobilenetv2 import MobileNetV2
from .resnet import ResNet, ResNet18, ResNet34, ResNet50, ResNet101, ResNet152
from .preact_resnet import (PreActResNet, PreActResNet18, PreActResNet34, PreActResNet50,
                            PreActResNet101, PreActResNet152)
from .vgg import VGG
from .googlenet import GoogLeNet
from .densenet import DenseNet, DenseNet121, DenseNet161, DenseNet169, DenseNet201, densenet_cifar
from .resnext import ResNeXt, ResNeXt29_2x64d, ResNeXt29_4x64d, ResNeXt29_8x64d, ResNeXt29_32x4d
from .dpn import DPN, DPN26, DPN92
from .senet import SENet, SENet18
from .shufflenet import ShuffleNet, ShuffleNetG2, ShuffleNetG3
from .shufflenetv2 import ShuffleNetV2
from .efficientnet import EfficientNet, EfficientNetB0
from .regnet import RegNet, RegNetX_200MF, RegNetX_400MF, RegNetY_400MF
from .pnasnet import PNASNet, PNASNetA, PNASNetB
from .dla_simple import SimpleDLA
from .dla import DLA

REGISTRY = {
    "LeNet": LeNet, "MobileNet": MobileNet, "MobileNetV2": MobileNetV2,
    "ResNet18": ResNet18, "ResNet34": ResNet34, "ResNet50": ResNet50, "ResNet101": ResNet101,
    "ResNet152": ResNet152,
    "PreActResNet18": PreActResNet18, "PreActResNet34": PreActResNet34,
    "PreActResNet50": PreActResNet50, "PreActResNet101": PreActResNet101,
    "PreActResNet152": PreActResNet152,
    "VGG11": lambda: VGG("VGG11"), "VGG13": lambda: VGG("VGG13"), "VGG16": lambda: VGG("VGG16"),
    "VGG19": lambda: VGG("VGG19"),
    "GoogLeNet": GoogLeNet,
    "DenseNet121": DenseNet121, "DenseNet169": DenseNet169, "DenseNet201": DenseNet201,
    "DenseNet161": DenseNet161, "densenet_cifar": densenet_cifar,
    "ResNeXt29_2x64d": ResNeXt29_2x64d, "ResNeXt29_4x64d": ResNeXt29_4x64d,
    "ResNeXt29_8x64d": ResNeXt29_8x64d, "ResNeXt29_32x4d": ResNeXt29_32x4d,
    "DPN26": DPN26, "DPN92": DPN92, "SENet18": SENet18,
    "ShuffleNetG2": ShuffleNetG2, "ShuffleNetG3": ShuffleNetG3,
    "ShuffleNetV2_0.5": lambda: ShuffleNetV2(0.5), "ShuffleNetV2_1": lambda: ShuffleNetV2(1),
    "ShuffleNetV2_1.5": lambda: ShuffleNetV2(1.5), "ShuffleNetV2_2": lambda: ShuffleNetV2(2),
    "EfficientNetB0": EfficientNetB0,
    "RegNetX_200MF": RegNetX_200MF, "RegNetX_400MF": RegNetX_400MF, "RegNetY_400MF": RegNetY_400MF,
    "PNASNetA": PNASNetA, "PNASNetB": PNASNetB, "SimpleDLA": SimpleDLA, "DLA": DLA,
}


def build(name: str):
    try:
        return REGISTRY[name]()
    except KeyError:
        raise KeyError(f"unknown model {name!r}; choose from {sorted(REGISTRY)}") from None
