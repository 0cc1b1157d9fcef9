"""VGG-11/16/19 (torchvision-compatible keys). VGG16 is one of the three
models of the reference's published scalability plots: 138 M
parameters, i.e. the communication-heavy end of the benchmark set
(/root/reference/srcs/python/kungfu/tensorflow/v1/benchmarks/model_sizes.py:1-5)."""
import torch
import torch.nn as nn

_CFG = {
    11: [64, 'M', 128, 'M', 256, 256, 'M', 512, 512, 'M', 512, 512, 'M'],
    16: [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 'M', 512, 512, 512, 'M', 512, 512, 512, 'M'],
    19: [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 256, 'M', 512, 512, 512, 512, 'M', 512, 512, 512, 512, 'M'],
}


class VGG(nn.Module):
    def __init__(self, depth=16, num_classes=1000, dropout=0.5, **_):
        super().__init__()
        layers, c = [], 3
        for v in _CFG[depth]:
            if v == 'M':
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [nn.Conv2d(c, v, 3, padding=1), nn.ReLU(inplace=True)]
                c = v
        self.features = nn.Sequential(*layers)
        self.avgpool = nn.AdaptiveAvgPool2d((7, 7))
        self.classifier = nn.Sequential(
            nn.Linear(512 * 7 * 7, 4096), nn.ReLU(True), nn.Dropout(dropout),
            nn.Linear(4096, 4096), nn.ReLU(True), nn.Dropout(dropout), nn.Linear(4096, num_classes))

    def forward(self, x):
        x = self.avgpool(self.features(x))
        return self.classifier(torch.flatten(x, 1))


def vgg11(**kw):
    """VGG-11 (torchvision keys)."""
    return VGG(11, **kw)


def vgg16(**kw):
    """VGG-16 (138 M parameters; the communication-heavy benchmark model)."""
    return VGG(16, **kw)


def vgg19(**kw):
    """VGG-19."""
    return VGG(19, **kw)
