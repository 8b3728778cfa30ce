"""Evaluate an MNIST checkpoint: ``python -m hetseq_b200.eval_mnist --model_ckpt F --mnist_dir D``
(reference: eval_mnist.py:40-100). Loads ``checkpoint['model']`` into ``MNISTNet``, runs the test
split at batch 64 and prints the average loss and accuracy. Device is a flag (the reference
hard-codes ``cuda``, eval_mnist.py:47); the test split is read from ``MNIST/processed/test.pt``
directly, so torchvision is not needed.
"""
from __future__ import annotations

import argparse
import os

import torch
import torch.nn.functional as F

from hetseq_b200.data.mnist_dataset import MNISTDataset
from hetseq_b200.models.mnist import MNISTNet
from hetseq_b200.utils.checkpoint_utils import load_checkpoint_to_cpu


def evaluate(model, dataset, device, batch_size=64):
    model.eval()
    test_loss, correct = 0.0, 0
    with torch.no_grad():
        for start in range(0, len(dataset), batch_size):
            data, target = dataset.fetch_batch(list(range(start, min(len(dataset), start + batch_size))))
            data, target = data.to(device), target.to(device)
            output = F.log_softmax(model.logits(data), dim=1)
            test_loss += F.nll_loss(output, target, reduction='sum').item()
            correct += output.argmax(dim=1).eq(target).sum().item()
    n = len(dataset)
    return test_loss / n, correct, n


def main(argv=None):
    parser = argparse.ArgumentParser(description='PyTorch MNIST Evaluation')
    parser.add_argument('--model_ckpt', '--model-ckpt', type=str, required=True, help='checkpoint file')
    parser.add_argument('--mnist_dir', '--mnist-dir', type=str, required=True, help='directory holding MNIST/processed')
    parser.add_argument('--device', type=str, default='cuda' if torch.cuda.is_available() else 'cpu')
    parser.add_argument('--batch-size', type=int, default=64)
    args = parser.parse_args(argv)

    path = args.mnist_dir
    for cand in (os.path.join(path, 'MNIST', 'processed', 'test.pt'), os.path.join(path, 'processed', 'test.pt'),
                 os.path.join(path, 'test.pt'), path):
        if os.path.isfile(cand):
            path = cand
            break
    dataset = MNISTDataset(path)
    model = MNISTNet()
    model.load_state_dict(load_checkpoint_to_cpu(args.model_ckpt)['model'])
    model.to(args.device)
    loss, correct, n = evaluate(model, dataset, args.device, args.batch_size)
    print('\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n'.format(loss, correct, n, 100. * correct / n))
    return loss, correct / n


if __name__ == '__main__':
    main()
