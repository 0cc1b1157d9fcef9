"""Model zoo used by the benchmarks and examples (all defined locally, random
init; ResNet / DenseNet / VGG keep torchvision's state_dict keys)."""
from .bert import BertConfig, BertForPreTraining, bert_base, bert_large, bert_tiny  # noqa: F401
from .densenet import DenseNet, densenet121, densenet169, densenet201  # noqa: F401
from .inception import Inception3, inception_v3  # noqa: F401
from .resnet import (ResNet, resnet18, resnet34, resnet50, resnet101,  # noqa: F401
                     resnet152, to_b200_layout)
from .simple import SLP, MnistCNN, mlp, slp  # noqa: F401
from .vgg import VGG, vgg11, vgg16, vgg19  # noqa: F401
