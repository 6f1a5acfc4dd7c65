"""Small workloads: the 2-conv/2-fc MNIST net of ``examples/torch/pytorch_mnist.py:86-102``
(21 840 parameters, 8 tensors) and the DAWNBench ResNet-9 of
``examples/dist/CIFAR10-dawndist/dawn.py:26-63``."""
import torch
from torch import nn
import torch.nn.functional as F


class MnistNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        x = F.relu(F.max_pool2d(self.drop(self.conv2(x)), 2))
        x = F.relu(self.fc1(x.flatten(1)))
        return F.log_softmax(self.fc2(F.dropout(x, training=self.training)), dim=1)


def _conv_bn(c_in, c_out):
    return nn.Sequential(nn.Conv2d(c_in, c_out, 3, padding=1, bias=False), nn.BatchNorm2d(c_out), nn.ReLU(True))


class _Residual(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.a, self.b = _conv_bn(c, c), _conv_bn(c, c)

    def forward(self, x):
        return x + self.b(self.a(x))


def resnet9(num_classes=10, width=64, scale=0.125):
    """prep -> (conv+pool, residual) -> conv+pool -> (conv+pool, residual) -> pool -> linear."""
    w = width

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.body = nn.Sequential(
                _conv_bn(3, w),
                _conv_bn(w, 2 * w), nn.MaxPool2d(2), _Residual(2 * w),
                _conv_bn(2 * w, 4 * w), nn.MaxPool2d(2),
                _conv_bn(4 * w, 8 * w), nn.MaxPool2d(2), _Residual(8 * w),
                nn.AdaptiveMaxPool2d(1), nn.Flatten())
            self.head = nn.Linear(8 * w, num_classes, bias=False)

        def forward(self, x):
            return self.head(self.body(x)) * scale

    return Net()
