"""Small models of the reference's examples and tests: the single-layer
perceptron of the MNIST np-invariance test
(/root/reference/tests/python/integration/test_mnist_slp.py) and the MNIST CNN of
/root/reference/examples/torch_mnist_example.py."""
import torch.nn as nn
import torch.nn.functional as F


class SLP(nn.Module):
    def __init__(self, input_size=784, num_classes=10, zero_init=True):
        super().__init__()
        self.fc = nn.Linear(input_size, num_classes)
        if zero_init:
            nn.init.zeros_(self.fc.weight)
            nn.init.zeros_(self.fc.bias)

    def forward(self, x):
        return self.fc(x.reshape(x.shape[0], -1))


class MnistCNN(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, 5)
        self.conv2 = nn.Conv2d(10, 20, 5)
        self.drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        x = F.relu(F.max_pool2d(self.drop(self.conv2(x)), 2))
        x = F.relu(self.fc1(x.flatten(1)))
        return F.log_softmax(self.fc2(F.dropout(x, training=self.training)), dim=1)


def slp(input_size=784, num_classes=10):
    """Single-layer perceptron (reference: v1/benchmarks/mnist.py slp)."""
    return SLP(input_size, num_classes, zero_init=False)


def mlp(input_size=784, hidden=(500, 300), num_classes=10):
    """Multi-layer perceptron with ReLU hidden layers (reference: v1/benchmarks/mnist.py mlp)."""
    layers, c = [nn.Flatten()], input_size
    for h in hidden:
        layers += [nn.Linear(c, h), nn.ReLU(inplace=True)]
        c = h
    layers.append(nn.Linear(c, num_classes))
    return nn.Sequential(*layers)
