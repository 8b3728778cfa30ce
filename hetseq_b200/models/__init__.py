"""Model families: BERT (pre-training + fine-tuning heads) and the MNIST CNN."""
