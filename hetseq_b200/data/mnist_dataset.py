"""MNIST in torchvision's ``processed/*.pt`` layout (reference: data/mnist_dataset.py).

A processed file is ``torch.save((uint8 [N,28,28], int64 [N]))``. ``__getitem__`` returns
``(float32 [1,28,28] normalised with mean 0.1307 / std 0.3081, int label)`` -- numerically the
reference's PIL -> ToTensor -> Normalize pipeline (data/mnist_dataset.py:16-19,47-52) without the
PIL / torchvision dependency. ``make_synthetic_mnist`` writes a small fake corpus in the same
layout for offline tests.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.utils.data
from torch.utils.data.dataloader import default_collate

MNIST_MEAN, MNIST_STD = 0.1307, 0.3081


class MNISTDataset(torch.utils.data.Dataset):
    def __init__(self, path):
        self.path = path
        self.read_data(path)

    def read_data(self, path):
        data = torch.load(path, weights_only=False)
        self.image, self.label = data[0], data[1]
        self._len = len(self.image)

    def __getitem__(self, index):
        img = self.image[index].to(torch.float32).div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD).unsqueeze(0)
        return img, int(self.label[index])

    def fetch_batch(self, indices):
        if len(indices) == 0:
            return None
        idx = torch.as_tensor(np.asarray(indices, dtype=np.int64))
        img = self.image[idx].to(torch.float32).div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD).unsqueeze(1)
        return [img, self.label[idx].to(torch.int64)]

    def __len__(self):
        return self._len

    def ordered_indices(self):
        return np.arange(len(self))

    def num_tokens(self, index: int):
        return 1

    def sizes_array(self):
        return np.ones(len(self), dtype=np.int64)

    def collater(self, samples):
        if len(samples) == 0:
            return None
        return default_collate(samples)

    def set_epoch(self, epoch):
        pass


def make_synthetic_mnist(root, n_train=512, n_test=128, seed=0):
    """Write ``root/MNIST/processed/{training,test}.pt`` with class-dependent blobs (learnable)."""
    g = torch.Generator().manual_seed(seed)
    out = os.path.join(root, 'MNIST', 'processed')
    os.makedirs(out, exist_ok=True)
    for name, n in (('training.pt', n_train), ('test.pt', n_test)):
        labels = torch.randint(0, 10, (n,), generator=g)
        img = torch.rand(n, 28, 28, generator=g) * 40
        for c in range(10):
            r0, c0 = 2 + 2 * c, 24 - 2 * c
            sel = labels == c
            img[sel, r0:r0 + 6, c0 - 4:c0 + 2] += 200
        torch.save((img.clamp(0, 255).to(torch.uint8), labels.to(torch.int64)), os.path.join(out, name))
    return out
