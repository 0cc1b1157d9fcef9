"""Inception v3 (torchvision-compatible module names and state_dict keys).

The third model of the reference's published scalability plots (BASELINE.md 2.1 / 2.2:
ResNet-50, VGG16, InceptionV3; the reference takes it from tf.keras.applications by name,
/root/reference/benchmarks/system/benchmark_kungfu.py). 299 x 299 input, every convolution is
conv + BatchNorm(eps=1e-3) + ReLU, which with ``fused_bn=True`` runs as cuDNN conv + ONE fused
BN+ReLU kernel (csrc/kernels/bn_act.cu) on channels-last bf16; all channel counts are multiples of 8.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .resnet import BNAct


class BasicConv2d(nn.Module):
    def __init__(self, cin, cout, fused=False, **kw):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, bias=False, **kw)
        self.bn = BNAct(cout, relu=True, fused=fused, eps=0.001)

    def forward(self, x):
        return self.bn(self.conv(x))


class InceptionA(nn.Module):
    def __init__(self, cin, pool_features, f):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 64, f, kernel_size=1)
        self.branch5x5_1 = BasicConv2d(cin, 48, f, kernel_size=1)
        self.branch5x5_2 = BasicConv2d(48, 64, f, kernel_size=5, padding=2)
        self.branch3x3dbl_1 = BasicConv2d(cin, 64, f, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, f, kernel_size=3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, f, kernel_size=3, padding=1)
        self.branch_pool = BasicConv2d(cin, pool_features, f, kernel_size=1)

    def forward(self, x):
        b1 = self.branch1x1(x)
        b5 = self.branch5x5_2(self.branch5x5_1(x))
        b3 = self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x)))
        bp = self.branch_pool(F.avg_pool2d(x, 3, stride=1, padding=1))
        return torch.cat([b1, b5, b3, bp], 1)


class InceptionB(nn.Module):
    def __init__(self, cin, f):
        super().__init__()
        self.branch3x3 = BasicConv2d(cin, 384, f, kernel_size=3, stride=2)
        self.branch3x3dbl_1 = BasicConv2d(cin, 64, f, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, f, kernel_size=3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, f, kernel_size=3, stride=2)

    def forward(self, x):
        b3 = self.branch3x3(x)
        bd = self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x)))
        return torch.cat([b3, bd, F.max_pool2d(x, 3, stride=2)], 1)


class InceptionC(nn.Module):
    def __init__(self, cin, c7, f):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 192, f, kernel_size=1)
        self.branch7x7_1 = BasicConv2d(cin, c7, f, kernel_size=1)
        self.branch7x7_2 = BasicConv2d(c7, c7, f, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7_3 = BasicConv2d(c7, 192, f, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_1 = BasicConv2d(cin, c7, f, kernel_size=1)
        self.branch7x7dbl_2 = BasicConv2d(c7, c7, f, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_3 = BasicConv2d(c7, c7, f, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7dbl_4 = BasicConv2d(c7, c7, f, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_5 = BasicConv2d(c7, 192, f, kernel_size=(1, 7), padding=(0, 3))
        self.branch_pool = BasicConv2d(cin, 192, f, kernel_size=1)

    def forward(self, x):
        b1 = self.branch1x1(x)
        b7 = self.branch7x7_3(self.branch7x7_2(self.branch7x7_1(x)))
        bd = self.branch7x7dbl_5(self.branch7x7dbl_4(self.branch7x7dbl_3(self.branch7x7dbl_2(self.branch7x7dbl_1(x)))))
        bp = self.branch_pool(F.avg_pool2d(x, 3, stride=1, padding=1))
        return torch.cat([b1, b7, bd, bp], 1)


class InceptionD(nn.Module):
    def __init__(self, cin, f):
        super().__init__()
        self.branch3x3_1 = BasicConv2d(cin, 192, f, kernel_size=1)
        self.branch3x3_2 = BasicConv2d(192, 320, f, kernel_size=3, stride=2)
        self.branch7x7x3_1 = BasicConv2d(cin, 192, f, kernel_size=1)
        self.branch7x7x3_2 = BasicConv2d(192, 192, f, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7x3_3 = BasicConv2d(192, 192, f, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7x3_4 = BasicConv2d(192, 192, f, kernel_size=3, stride=2)

    def forward(self, x):
        b3 = self.branch3x3_2(self.branch3x3_1(x))
        b7 = self.branch7x7x3_4(self.branch7x7x3_3(self.branch7x7x3_2(self.branch7x7x3_1(x))))
        return torch.cat([b3, b7, F.max_pool2d(x, 3, stride=2)], 1)


class InceptionE(nn.Module):
    def __init__(self, cin, f):
        super().__init__()
        self.branch1x1 = BasicConv2d(cin, 320, f, kernel_size=1)
        self.branch3x3_1 = BasicConv2d(cin, 384, f, kernel_size=1)
        self.branch3x3_2a = BasicConv2d(384, 384, f, kernel_size=(1, 3), padding=(0, 1))
        self.branch3x3_2b = BasicConv2d(384, 384, f, kernel_size=(3, 1), padding=(1, 0))
        self.branch3x3dbl_1 = BasicConv2d(cin, 448, f, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(448, 384, f, kernel_size=3, padding=1)
        self.branch3x3dbl_3a = BasicConv2d(384, 384, f, kernel_size=(1, 3), padding=(0, 1))
        self.branch3x3dbl_3b = BasicConv2d(384, 384, f, kernel_size=(3, 1), padding=(1, 0))
        self.branch_pool = BasicConv2d(cin, 192, f, kernel_size=1)

    def forward(self, x):
        b1 = self.branch1x1(x)
        b3 = self.branch3x3_1(x)
        b3 = torch.cat([self.branch3x3_2a(b3), self.branch3x3_2b(b3)], 1)
        bd = self.branch3x3dbl_2(self.branch3x3dbl_1(x))
        bd = torch.cat([self.branch3x3dbl_3a(bd), self.branch3x3dbl_3b(bd)], 1)
        bp = self.branch_pool(F.avg_pool2d(x, 3, stride=1, padding=1))
        return torch.cat([b1, b3, bd, bp], 1)


class InceptionAux(nn.Module):
    def __init__(self, cin, num_classes, f):
        super().__init__()
        self.conv0 = BasicConv2d(cin, 128, f, kernel_size=1)
        self.conv1 = BasicConv2d(128, 768, f, kernel_size=5)
        self.fc = nn.Linear(768, num_classes)

    def forward(self, x):
        x = F.avg_pool2d(x, 5, stride=3)
        x = self.conv1(self.conv0(x))
        x = F.adaptive_avg_pool2d(x, (1, 1))
        return self.fc(torch.flatten(x, 1))


class Inception3(nn.Module):
    def __init__(self, num_classes=1000, aux_logits=False, dropout=0.5, fused_bn=False, **_):
        super().__init__()
        f = fused_bn
        self.aux_logits = aux_logits
        self.Conv2d_1a_3x3 = BasicConv2d(3, 32, f, kernel_size=3, stride=2)
        self.Conv2d_2a_3x3 = BasicConv2d(32, 32, f, kernel_size=3)
        self.Conv2d_2b_3x3 = BasicConv2d(32, 64, f, kernel_size=3, padding=1)
        self.Conv2d_3b_1x1 = BasicConv2d(64, 80, f, kernel_size=1)
        self.Conv2d_4a_3x3 = BasicConv2d(80, 192, f, kernel_size=3)
        self.Mixed_5b = InceptionA(192, 32, f)
        self.Mixed_5c = InceptionA(256, 64, f)
        self.Mixed_5d = InceptionA(288, 64, f)
        self.Mixed_6a = InceptionB(288, f)
        self.Mixed_6b = InceptionC(768, 128, f)
        self.Mixed_6c = InceptionC(768, 160, f)
        self.Mixed_6d = InceptionC(768, 160, f)
        self.Mixed_6e = InceptionC(768, 192, f)
        if aux_logits:
            self.AuxLogits = InceptionAux(768, num_classes, f)
        self.Mixed_7a = InceptionD(768, f)
        self.Mixed_7b = InceptionE(1280, f)
        self.Mixed_7c = InceptionE(2048, f)
        self.dropout = nn.Dropout(dropout)
        self.fc = nn.Linear(2048, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                nn.init.trunc_normal_(m.weight, mean=0.0, std=0.1, a=-2, b=2)

    def forward(self, x):
        x = self.Conv2d_2b_3x3(self.Conv2d_2a_3x3(self.Conv2d_1a_3x3(x)))
        x = F.max_pool2d(x, 3, stride=2)
        x = self.Conv2d_4a_3x3(self.Conv2d_3b_1x1(x))
        x = F.max_pool2d(x, 3, stride=2)
        x = self.Mixed_5d(self.Mixed_5c(self.Mixed_5b(x)))
        x = self.Mixed_6e(self.Mixed_6d(self.Mixed_6c(self.Mixed_6b(self.Mixed_6a(x)))))
        aux = self.AuxLogits(x) if (self.aux_logits and self.training) else None
        x = self.Mixed_7c(self.Mixed_7b(self.Mixed_7a(x)))
        x = torch.flatten(F.adaptive_avg_pool2d(x, (1, 1)), 1)
        x = self.fc(self.dropout(x))
        return (x, aux) if aux is not None else x


def inception_v3(**kw):
    """Inception v3 for 299x299 inputs (23.8 M parameters, 27.2 M with aux_logits=True). Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), aux_logits, dropout."""
    return Inception3(**kw)
