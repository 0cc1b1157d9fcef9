"""ResNet family (18/34/50/101/152), torchvision-compatible state_dict keys.

The reference benchmarks take models from tf.keras.applications /
torchvision.models by name (/root/reference/benchmarks/system/
benchmark_kungfu_torch.py:61, benchmark_kungfu.py); this is the same
architecture defined locally so the benchmark has no external dependency, laid
out for B200: channels-last activations, bf16 convolutions, fp32 batch-norm
statistics, and (optionally) the fused BN+ReLU(+residual) kernels of
csrc/kernels/bn_act.cu instead of separate normalisation / activation / add
passes over HBM.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class BNAct(nn.BatchNorm2d):
    """BatchNorm2d (+ optional residual add) (+ optional ReLU).

    With ``fused=True`` on CUDA channels-last bf16/fp16 input this runs the
    hand-written kernels (kungfu_b200.ops.bn_act); otherwise plain PyTorch ops.
    Parameters/buffers are those of nn.BatchNorm2d, so checkpoints interchange.
    """

    def __init__(self, num_features, relu=True, fused=False, **kw):
        super().__init__(num_features, **kw)
        self.relu = relu
        self.fused = fused

    def forward(self, x, residual=None):
        if self.fused and x.is_cuda and x.dtype in (torch.bfloat16, torch.float16):
            from ..ops.bn_act import bn_act
            return bn_act(x, residual, self.weight, self.bias, self.running_mean,
                          self.running_var, self.training, self.momentum, self.eps,
                          self.relu, self.num_batches_tracked)
        y = super().forward(x)
        if residual is not None:
            y = y + residual
        return F.relu(y, inplace=True) if self.relu else y


def conv3x3(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, 3, stride=stride, padding=1, bias=False)


def conv1x1(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, 1, stride=stride, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride=1, downsample=None, fused=False):
        super().__init__()
        self.conv1 = conv3x3(cin, planes, stride)
        self.bn1 = BNAct(planes, relu=True, fused=fused)
        self.conv2 = conv3x3(planes, planes)
        self.bn2 = BNAct(planes, relu=True, fused=fused)
        self.downsample = downsample

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        return self.bn2(self.conv2(out), identity)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride=1, downsample=None, fused=False):
        super().__init__()
        self.conv1 = conv1x1(cin, planes)
        self.bn1 = BNAct(planes, relu=True, fused=fused)
        self.conv2 = conv3x3(planes, planes, stride)
        self.bn2 = BNAct(planes, relu=True, fused=fused)
        self.conv3 = conv1x1(planes, planes * 4)
        self.bn3 = BNAct(planes * 4, relu=True, fused=fused)
        self.downsample = downsample

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        return self.bn3(self.conv3(out), identity)


class _Downsample(nn.Sequential):
    """conv1x1 + BN, keys '0' and '1' like torchvision's nn.Sequential."""

    def __init__(self, cin, cout, stride, fused):
        super().__init__(conv1x1(cin, cout, stride), BNAct(cout, relu=False, fused=fused))


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000, fused_bn=False, zero_init_residual=False, s2d_stem=False):
        super().__init__()
        self.inplanes = 64
        self.fused_bn = fused_bn
        self.s2d_stem = s2d_stem
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bn1 = BNAct(64, relu=True, fused=fused_bn)
        from ..ops.pool import MaxPool3x3s2
        self.maxpool = MaxPool3x3s2(fused=fused_bn)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, Bottleneck):
                    nn.init.constant_(m.bn3.weight, 0)
                elif isinstance(m, BasicBlock):
                    nn.init.constant_(m.bn2.weight, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = _Downsample(self.inplanes, planes * block.expansion, stride, self.fused_bn)
        layers = [block(self.inplanes, planes, stride, downsample, self.fused_bn)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes, fused=self.fused_bn))
        return nn.Sequential(*layers)

    def _stem(self, x):
        """conv1 (7x7, stride 2, 3 input channels). With ``s2d_stem`` the same
        convolution is evaluated as a 4x4 stride-1 convolution over a 2x2
        space-to-depth view of the padded input (12 channels): identical maths
        and the same [64,3,7,7] parameter, but a shape tensor cores like far
        better than C_in = 3."""
        if not self.s2d_stem or x.shape[2] % 2 or x.shape[3] % 2:
            return self.conv1(x)
        w = self.conv1.weight
        N, C, H, W = x.shape
        Hh, Wh = (H + 6) // 2, (W + 6) // 2
        xs = F.pad(x, (3, 3, 3, 3)).view(N, C, Hh, 2, Wh, 2).permute(0, 3, 5, 1, 2, 4).reshape(N, 4 * C, Hh, Wh)
        ws = F.pad(w, (0, 1, 0, 1)).view(w.shape[0], C, 4, 2, 4, 2).permute(0, 3, 5, 1, 2, 4).reshape(w.shape[0], 4 * C, 4, 4)
        return F.conv2d(xs.contiguous(memory_format=torch.channels_last), ws.contiguous(memory_format=torch.channels_last))

    def forward(self, x):
        x = self.maxpool(self.bn1(self._stem(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        if self.fused_bn:
            from ..ops.pool import global_avg_pool
            x = global_avg_pool(x)
        else:
            x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x)


def resnet18(**kw):
    """ResNet-18 (BasicBlock, [2, 2, 2, 2]). Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), s2d_stem."""
    return ResNet(BasicBlock, [2, 2, 2, 2], **kw)


def resnet34(**kw):
    """ResNet-34 (BasicBlock, [3, 4, 6, 3]). Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), s2d_stem."""
    return ResNet(BasicBlock, [3, 4, 6, 3], **kw)


def resnet50(**kw):
    """ResNet-50 (Bottleneck, [3, 4, 6, 3]; 25.6 M parameters): the headline benchmark model. Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), s2d_stem."""
    return ResNet(Bottleneck, [3, 4, 6, 3], **kw)


def resnet101(**kw):
    """ResNet-101 (Bottleneck, [3, 4, 23, 3]). Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), s2d_stem."""
    return ResNet(Bottleneck, [3, 4, 23, 3], **kw)


def resnet152(**kw):
    """ResNet-152 (Bottleneck, [3, 8, 36, 3]). Keyword arguments: num_classes, fused_bn (fused BN+ReLU(+residual) kernels on channels-last bf16), s2d_stem."""
    return ResNet(Bottleneck, [3, 8, 36, 3], **kw)


def to_b200_layout(model, dtype=torch.bfloat16):
    """Channels-last, `dtype` weights for conv/linear, fp32 batch-norm (the
    O2-style mixed precision the fused optimizers keep fp32 masters for)."""
    model = model.to(memory_format=torch.channels_last)
    for m in model.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            m.to(dtype)
        elif isinstance(m, nn.modules.batchnorm._BatchNorm):
            m.float()
    return model
