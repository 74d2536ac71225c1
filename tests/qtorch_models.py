"""Small eager-mode-quantized torch models for the importer tests (tachikoma_b200/frontend.py): built, calibrated and
converted on the CPU with torch.ao.quantization, then traced.  Not collected by pytest (no test_ prefix)."""
import warnings

import numpy as np
import torch
import torch.nn as nn

warnings.filterwarnings("ignore")
import torch.ao.quantization as tq  # noqa: E402


class Block(nn.Module):
    """ResNet BasicBlock with a FloatFunctional residual add (torchvision.models.quantization.resnet layout)."""

    def __init__(self, c, stride=1, cout=None):
        super().__init__()
        cout = cout or c
        self.c1 = nn.Conv2d(c, cout, 3, stride, 1, bias=False)
        self.b1 = nn.BatchNorm2d(cout)
        self.r1 = nn.ReLU()
        self.c2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.b2 = nn.BatchNorm2d(cout)
        self.down = None
        if stride != 1 or cout != c:
            self.down = nn.Sequential(nn.Conv2d(c, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))
        self.ff = nn.quantized.FloatFunctional()

    def forward(self, x):
        y = self.b2(self.c2(self.r1(self.b1(self.c1(x)))))
        return self.ff.add_relu(y, x if self.down is None else self.down(x))


class TinyResNet(nn.Module):
    def __init__(self, width=16, classes=10):
        super().__init__()
        self.q, self.dq = tq.QuantStub(), tq.DeQuantStub()
        self.c = nn.Conv2d(3, width, 7, 2, 3, bias=False)
        self.b = nn.BatchNorm2d(width)
        self.r = nn.ReLU()
        self.mp = nn.MaxPool2d(3, 2, 1)
        self.l1 = Block(width)
        self.l2 = Block(width, 2, 2 * width)
        self.ap = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(2 * width, classes)

    def forward(self, x):
        x = self.mp(self.r(self.b(self.c(self.q(x)))))
        x = self.ap(self.l2(self.l1(x)))
        return self.dq(self.fc(torch.flatten(x, 1)))

    def fuse(self):
        tq.fuse_modules(self, [["c", "b", "r"]], inplace=True)
        for blk in (self.l1, self.l2):
            tq.fuse_modules(blk, [["c1", "b1", "r1"], ["c2", "b2"]], inplace=True)
            if blk.down is not None:
                tq.fuse_modules(blk.down, [["0", "1"]], inplace=True)


class CatMulNet(nn.Module):
    """conv (with bias, no relu) -> two branches (1x1 conv_relu, depthwise 3x3) -> quantized::cat -> quantized::mul ->
    avg_pool2d."""

    def __init__(self):
        super().__init__()
        self.q, self.dq = tq.QuantStub(), tq.DeQuantStub()
        self.c0 = nn.Conv2d(4, 8, 3, 1, 1)
        self.ca = nn.Conv2d(8, 8, 1)
        self.ra = nn.ReLU()
        self.cb = nn.Conv2d(8, 8, 3, 1, 1, groups=8)
        self.ff = nn.quantized.FloatFunctional()
        self.ff2 = nn.quantized.FloatFunctional()
        self.pool = nn.AvgPool2d(2, 2)

    def forward(self, x):
        x = self.c0(self.q(x))
        a, b = self.ra(self.ca(x)), self.cb(x)
        y = self.ff.cat([a, b], 1)
        return self.dq(self.pool(self.ff2.mul(y, y)))

    def fuse(self):
        tq.fuse_modules(self, [["ca", "ra"]], inplace=True)


class MLP(nn.Module):
    """quantized::linear_relu -> quantized::linear."""

    def __init__(self):
        super().__init__()
        self.q, self.dq = tq.QuantStub(), tq.DeQuantStub()
        self.f1 = nn.Linear(32, 24)
        self.r1 = nn.ReLU()
        self.f2 = nn.Linear(24, 6)

    def forward(self, x):
        return self.dq(self.f2(self.r1(self.f1(self.q(x)))))

    def fuse(self):
        tq.fuse_modules(self, [["f1", "r1"]], inplace=True)


class ActNet(nn.Module):
    """conv -> quantized::leaky_relu -> conv -> quantized::hardswish -> conv -> aten::sigmoid -> relu6 -> conv."""

    def __init__(self):
        super().__init__()
        self.q, self.dq = tq.QuantStub(), tq.DeQuantStub()
        self.c1 = nn.Conv2d(3, 8, 3, 1, 1)
        self.a1 = nn.LeakyReLU(0.1)
        self.c2 = nn.Conv2d(8, 8, 3, 1, 1)
        self.a2 = nn.Hardswish()
        self.c3 = nn.Conv2d(8, 8, 1)
        self.a3 = nn.Sigmoid()
        self.c4 = nn.Conv2d(8, 8, 1)
        self.a4 = nn.ReLU6()
        self.c5 = nn.Conv2d(8, 4, 1)

    def forward(self, x):
        x = self.a1(self.c1(self.q(x)))
        x = self.a2(self.c2(x))
        x = self.a3(self.c3(x))
        x = self.a4(self.c4(x))
        return self.dq(self.c5(x))

    def fuse(self):
        pass


def build(cls, shape, seed=0, per_channel=True):
    """-> (traced ScriptModule, eager quantized module).  BatchNorm statistics are randomized so that folding matters."""
    torch.manual_seed(seed)
    m = cls().eval()
    for mod in m.modules():
        if isinstance(mod, nn.BatchNorm2d):
            mod.running_mean.uniform_(-0.5, 0.5)
            mod.running_var.uniform_(0.5, 2.0)
            mod.weight.data.uniform_(0.5, 1.5)
            mod.bias.data.uniform_(-0.5, 0.5)
    m.fuse()
    torch.backends.quantized.engine = "fbgemm"
    m.qconfig = tq.get_default_qconfig("fbgemm") if per_channel else tq.default_qconfig
    tq.prepare(m, inplace=True)
    g = torch.Generator().manual_seed(seed + 1)
    for _ in range(4):
        m(torch.randn(shape, generator=g))
    tq.convert(m, inplace=True)
    x = torch.randn(shape, generator=g)
    with torch.no_grad():
        ts = torch.jit.trace(m, x)
    return ts, m


def sample(shape, seed=7):
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed)).numpy().astype(np.float32)
