"""Name -> class registries for tasks, optimizers and LR schedulers.

The reference selects these with hard-coded if/elif chains (train.py:43-46,
controller.py:111-124, options.py:61-85,166-202). Here a component registers itself with a
decorator and contributes its own flags through ``add_args(parser)``; adding a task therefore
needs no edit to the trainer or the option parser.
"""
from __future__ import annotations

from typing import Callable, Dict, Type

TASKS: Dict[str, type] = {}
OPTIMIZERS: Dict[str, type] = {}
LR_SCHEDULERS: Dict[str, type] = {}

_KINDS = {"task": TASKS, "optimizer": OPTIMIZERS, "lr_scheduler": LR_SCHEDULERS}


def _register(kind: str, name: str) -> Callable[[Type], Type]:
    table = _KINDS[kind]

    def deco(cls):
        key = name.lower()
        if key in table and table[key] is not cls:
            raise ValueError(f"{kind} '{name}' is already registered ({table[key]!r})")
        table[key] = cls
        cls.registered_name = name
        return cls

    return deco


def register_task(name: str):
    return _register("task", name)


def register_optimizer(name: str):
    return _register("optimizer", name)


def register_lr_scheduler(name: str):
    return _register("lr_scheduler", name)


def lookup(kind: str, name: str):
    table = _KINDS[kind]
    key = name.lower()
    if key not in table:
        raise ValueError(f"unsupported {kind}: {name!r} (known: {sorted(table)})")
    return table[key]


def ensure_builtin_components_loaded() -> None:
    """Import the modules whose decorators populate the registries."""
    import hetseq_b200.tasks  # noqa: F401
    import hetseq_b200.optim  # noqa: F401
