"""DenseNet-121/169/201 with torchvision-compatible state_dict keys.

BASELINE.json config 3 trains DenseNet-121 with the SMA optimizer; the
reference takes the model from tf.keras.applications by name
(/root/reference/benchmarks/system/benchmark_kungfu.py). Pre-activation
BN -> ReLU -> conv units use the fused BN+ReLU kernel when ``fused_bn``.
"""
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

from .resnet import BNAct


class _DenseLayer(nn.Module):
    def __init__(self, cin, growth, bn_size, fused):
        super().__init__()
        self.norm1 = BNAct(cin, relu=True, fused=fused)
        self.conv1 = nn.Conv2d(cin, bn_size * growth, 1, bias=False)
        self.norm2 = BNAct(bn_size * growth, relu=True, fused=fused)
        self.conv2 = nn.Conv2d(bn_size * growth, growth, 3, padding=1, bias=False)

    def forward(self, feats):
        x = torch.cat(feats, 1) if isinstance(feats, (list, tuple)) else feats
        x = self.conv1(self.norm1(x))
        return self.conv2(self.norm2(x))


class _DenseBlock(nn.ModuleDict):
    def __init__(self, n, cin, growth, bn_size, fused):
        super().__init__()
        for i in range(n):
            self['denselayer%d' % (i + 1)] = _DenseLayer(cin + i * growth, growth, bn_size, fused)

    def forward(self, x):
        feats = [x]
        for layer in self.values():
            feats.append(layer(feats))
        return torch.cat(feats, 1)


class _Transition(nn.Sequential):
    def __init__(self, cin, cout, fused):
        super().__init__(OrderedDict([
            ('norm', BNAct(cin, relu=True, fused=fused)),
            ('conv', nn.Conv2d(cin, cout, 1, bias=False)),
            ('pool', nn.AvgPool2d(2, 2)),
        ]))


class DenseNet(nn.Module):
    def __init__(self, growth=32, blocks=(6, 12, 24, 16), init_features=64, bn_size=4, num_classes=1000,
                 fused_bn=False):
        super().__init__()
        from ..ops.pool import MaxPool3x3s2
        self.fused_bn = fused_bn
        feats = OrderedDict([
            ('conv0', nn.Conv2d(3, init_features, 7, stride=2, padding=3, bias=False)),
            ('norm0', BNAct(init_features, relu=True, fused=fused_bn)),
            ('pool0', MaxPool3x3s2(fused=fused_bn)),
        ])
        c = init_features
        for i, n in enumerate(blocks):
            feats['denseblock%d' % (i + 1)] = _DenseBlock(n, c, growth, bn_size, fused_bn)
            c += n * growth
            if i != len(blocks) - 1:
                feats['transition%d' % (i + 1)] = _Transition(c, c // 2, fused_bn)
                c //= 2
        feats['norm5'] = BNAct(c, relu=True, fused=fused_bn)
        self.features = nn.Sequential(feats)
        self.classifier = nn.Linear(c, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight)
            elif isinstance(m, nn.Linear):
                nn.init.constant_(m.bias, 0)

    def forward(self, x):
        x = self.features(x)
        if self.fused_bn:
            from ..ops.pool import global_avg_pool
            x = global_avg_pool(x)
        else:
            x = torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)
        return self.classifier(x)


def densenet121(**kw):
    """DenseNet-121. Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16)."""
    return DenseNet(32, (6, 12, 24, 16), 64, **kw)


def densenet169(**kw):
    """DenseNet-169. Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16)."""
    return DenseNet(32, (6, 12, 32, 32), 64, **kw)


def densenet201(**kw):
    """DenseNet-201. Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16)."""
    return DenseNet(32, (6, 12, 48, 32), 64, **kw)
