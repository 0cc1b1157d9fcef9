"""Model zoo used by the benchmarks (all defined locally, random init)."""
from .resnet import (ResNet, resnet18, resnet34, resnet50, resnet101,  # noqa: F401
                     resnet152, to_b200_layout)
