#!/bin/bash
# Installs the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box).
#   1. pip builds what the reference's setup.py declares: the `data` package + the Cython extension
#      (--no-deps: the wheelhouse has no numpy/torch wheels; both are already in the image).
#   2. setup.py lists no py_modules, so the reference's flat top-level modules (train.py, controller.py,
#      bert_modeling.py, ...) are not part of the wheel; they are copied verbatim next to `data/`.
set -euo pipefail
cd "$(dirname "$0")/.."
REF=${REF:-/root/reference}
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target baseline/_ref --upgrade "$REF"
for f in "$REF"/*.py; do
  case "$(basename "$f")" in setup.py|__init__.py) continue;; esac
  cp "$f" baseline/_ref/
done
ls baseline/_ref
