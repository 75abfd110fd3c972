"""Plain PyTorch definitions of the models the benchmark configs name (the reference resolves
these names through the external cyy_torch_vision package)."""
import torch
from torch import nn


class LeNet5(nn.Module):
    def __init__(self, in_channels=1, num_classes=10, input_size=28) -> None:
        super().__init__()
        self.net = nn.Sequential(
            nn.Conv2d(in_channels, 6, 5, padding=2 if input_size == 28 else 0), nn.ReLU(), nn.MaxPool2d(2),
            nn.Conv2d(6, 16, 5), nn.ReLU(), nn.MaxPool2d(2), nn.Flatten(),
            nn.Linear(400, 120), nn.ReLU(), nn.Linear(120, 84), nn.ReLU(), nn.Linear(84, num_classes))

    def forward(self, x):
        return self.net(x)


class BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.shortcut = None
        if stride != 1 or cin != cout:
            self.shortcut = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        out = torch.relu(self.bn1(self.conv1(x)))
        out = self.bn2(self.conv2(out))
        return torch.relu(out + (x if self.shortcut is None else self.shortcut(x)))


class ResNet18(nn.Module):
    def __init__(self, in_channels=3, num_classes=10) -> None:
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(in_channels, 64, 3, 1, 1, bias=False), nn.BatchNorm2d(64), nn.ReLU())
        layers, cin = [], 64
        for cout, stride in [(64, 1), (64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1)]:
            layers.append(BasicBlock(cin, cout, stride))
            cin = cout
        self.layers = nn.Sequential(*layers)
        self.fc = nn.Linear(512, num_classes)

    def forward(self, x):
        x = self.layers(self.stem(x))
        return self.fc(torch.flatten(nn.functional.adaptive_avg_pool2d(x, 1), 1))


class DenseLayer(nn.Module):
    def __init__(self, cin, growth) -> None:
        super().__init__()
        self.norm, self.relu = nn.BatchNorm2d(cin), nn.ReLU()
        self.conv = nn.Conv2d(cin, growth, 3, padding=1, bias=False)

    def forward(self, x):
        return torch.cat([x, self.conv(self.relu(self.norm(x)))], 1)


class DenseNet40(nn.Module):
    def __init__(self, in_channels=3, num_classes=10, growth=12) -> None:
        super().__init__()
        ch = 2 * growth
        mods = [nn.Conv2d(in_channels, ch, 3, padding=1, bias=False)]
        for stage in range(3):
            for _ in range(12):
                mods.append(DenseLayer(ch, growth))
                ch += growth
            if stage != 2:
                mods += [nn.BatchNorm2d(ch), nn.ReLU(), nn.Conv2d(ch, ch, 1, bias=False), nn.AvgPool2d(2)]
        mods += [nn.BatchNorm2d(ch), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(ch, num_classes)]
        self.net = nn.Sequential(*mods)

    def forward(self, x):
        return self.net(x)


def create_model(name: str, spec: dict) -> nn.Module:
    c, classes = spec["shape"][0], spec["classes"]
    key = name.lower()
    if key == "lenet5":
        return LeNet5(c, classes, spec["shape"][1])
    if key == "resnet18":
        return ResNet18(c, classes)
    if key == "densenet40":
        return DenseNet40(c, classes)
    raise KeyError(f"model {name} is not available in the reference shims")
