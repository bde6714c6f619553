"""LeNet-5 style CNN for the MNIST examples (reference: ``scaletorch/models/lenet.py``)."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class LeNet(nn.Module):
    """Small CNN for 28 x 28 MNIST digits (two 3x3 convolutions, max-pool, two linear layers): the model of the DDP / FSDP examples."""
    def __init__(self, num_classes: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout(0.25)
        self.dropout2 = nn.Dropout(0.5)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, num_classes)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.conv1(x))
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = torch.flatten(self.dropout1(x), 1)
        x = self.dropout2(F.relu(self.fc1(x)))
        return F.log_softmax(self.fc2(x), dim=1)
