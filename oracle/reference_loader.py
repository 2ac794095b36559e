"""Import the live reference (`/root/reference/perceiver/{model,adapter,utils}.py`).

TEST INFRASTRUCTURE. Works only in the build container: the reference tree does not
travel to the GPU box, so nothing under `-m gpu`, smoke() or bench.py calls this.

`perceiver/__init__.py:2-6` of the reference eagerly imports `perceiver.lightning`
(pytorch_lightning / torchmetrics — not installed, not installable offline).  We register
a stub package object whose `__path__` points at the reference directory, so that
`perceiver.model`, `perceiver.adapter`, `perceiver.utils`, `perceiver.tokenizer` resolve to
the reference's own, unmodified files without executing its `__init__`.

The stub is registered under the name `perceiver` only for the duration of the import and
the loaded modules are returned; callers that also use this repo's drop-in `perceiver`
package must do so in another process.
"""
import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("PIO_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "perceiver", "model.py"))


def load():
    """Returns (model_module, adapter_module) of the unmodified reference."""
    if not available():
        raise RuntimeError(f"reference tree not present at {REFERENCE_ROOT}")
    for name in [n for n in sys.modules if n == "perceiver" or n.startswith("perceiver.")]:
        mod = sys.modules[name]
        path = getattr(mod, "__file__", None) or ""
        if not path.startswith(REFERENCE_ROOT) and not isinstance(getattr(mod, "__path__", None), list):
            raise RuntimeError("a different `perceiver` package is already imported in this process")
    if "perceiver" not in sys.modules:
        stub = types.ModuleType("perceiver")
        stub.__path__ = [os.path.join(REFERENCE_ROOT, "perceiver")]
        sys.modules["perceiver"] = stub
    model = importlib.import_module("perceiver.model")
    adapter = importlib.import_module("perceiver.adapter")
    return model, adapter
