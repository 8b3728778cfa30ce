"""hetseq_b200 -- a B200-native synchronous data-parallel training framework.

Capabilities follow BismarckBamfo/hetseq (flags, tasks, resumable iterators, fairseq-style
checkpoints); the architecture does not: one process per GPU, flat parameter/gradient buffers,
hand-written sm_100a kernels (csrc/) for the BERT hot ops and for the fused
reduce-scatter -> clip/unscale -> Adam -> all-gather gradient path over NVLink peer memory.
"""

__version__ = "0.1.0"

from hetseq_b200 import registry  # noqa: F401
