"""Tasks: dataset loading, batching glue, model construction and the per-batch train step."""
from hetseq_b200.tasks.base import Task  # noqa: F401
from hetseq_b200.tasks.bert import LanguageModelingTask, BertFineTuningTask  # noqa: F401
from hetseq_b200.tasks.mnist import MNISTTask  # noqa: F401
