"""MNIST CNN (reference: tasks.py:337-362, duplicated in eval_mnist.py:9-37).

conv(1->32,3) -> ReLU -> conv(32->64,3) -> ReLU -> maxpool2 -> Dropout2d(.25) -> flatten(9216)
-> fc(128) -> ReLU -> Dropout(.5) -> fc(10) -> log_softmax -> nll_loss. 1,199,882 parameters.
``forward(x, target)`` returns the loss (the framework's model contract); ``eval=True`` returns
``(output, loss)`` explicitly -- the reference's ``return loss if not eval else output, loss``
(eval_mnist.py:36) returns a tuple in both cases by operator precedence.
The convolutions stay on cuDNN / CPU kernels (config 1 is CPU plumbing).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class MNISTNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout2d(0.25)
        self.dropout2 = nn.Dropout(0.5)  # the reference applies Dropout2d to a 2-D tensor; same effect
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def logits(self, x):
        x = F.relu(self.conv1(x))
        x = F.relu(self.conv2(x))
        x = F.max_pool2d(x, 2)
        x = self.dropout1(x)
        x = torch.flatten(x, 1)
        x = F.relu(self.fc1(x))
        x = self.dropout2(x)
        return self.fc2(x)

    def forward(self, x, target, eval=False):
        output = F.log_softmax(self.logits(x), dim=1)
        loss = F.nll_loss(output, target.long())
        if eval:
            return output, loss
        return loss
