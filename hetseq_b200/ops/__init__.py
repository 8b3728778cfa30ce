"""Hot ops: Python bindings of the sm_100a kernel library plus their PyTorch restatements."""
