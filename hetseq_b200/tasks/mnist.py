"""MNIST task (reference: tasks.py:288-334) -- the CPU plumbing demo of BASELINE.json config 1."""
from __future__ import annotations

import os

from hetseq_b200.data.mnist_dataset import MNISTDataset
from hetseq_b200.registry import register_task
from hetseq_b200.tasks.base import Task


@register_task('mnist')
class MNISTTask(Task):
    @staticmethod
    def add_args(parser):
        parser.add_argument('--data', type=str, help='path including data')

    def build_model(self, args):
        from hetseq_b200.models.mnist import MNISTNet
        return MNISTNet()

    def load_dataset(self, split, **kwargs):
        path = self.args.data
        if path is None or not os.path.exists(path):
            raise FileNotFoundError('Dataset not found: ({})'.format(path))
        if os.path.isdir(path):
            if os.path.exists(os.path.join(path, 'MNIST/processed/')):
                path = os.path.join(path, 'MNIST/processed/')
            elif os.path.basename(os.path.normpath(path)) != 'processed':
                # the reference downloads through torchvision here (tasks.py:317-322)
                try:
                    from torchvision import datasets
                    datasets.MNIST(path, train=True, download=True)
                except Exception as e:
                    raise FileNotFoundError(
                        'no MNIST/processed/ under {} and the torchvision download failed ({})'.format(path, e))
                path = os.path.join(path, 'MNIST/processed/')
        files = [os.path.join(path, f) for f in os.listdir(path)] if os.path.isdir(path) else [path]
        files = sorted([f for f in files if split in os.path.basename(f)])
        assert len(files) == 1, 'no suitable file in split ***{}***'.format(split)
        dataset = MNISTDataset(files[0])
        print('| loaded {} sentences from: {}'.format(len(dataset), path), flush=True)
        self.datasets[split] = dataset
        print('| loading finished')
