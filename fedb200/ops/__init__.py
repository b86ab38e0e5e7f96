"""Op wrappers over the hand-written sm_100a kernels in ``csrc/`` (loaded by :mod:`._lib`).

``cuda_ops`` is imported lazily so CPU-only runs never touch the shared library; on a machine
with a GPU a missing ``libfedb200.so`` raises immediately (no silent PyTorch fallback)."""
import importlib


def __getattr__(name):
    if name == "cuda_ops":
        return importlib.import_module(".cuda_ops", __name__)
    raise AttributeError(name)
