"""gaussian-splatterer_b200: B200-native (sm_100a) training hot path of osreboot/Gaussian-Splatterer.

The product is libgsb200.so (hand-written CUDA kernels + the C ABI of include/gsb.h) and the C++
headers in include/compat/ that give it the reference's class signatures.  This Python package is
the harness around it: a ctypes binding, mirrors of the reference's Rasterizer / Trainer /
ModelSplats* / Camera / Project interfaces for tests and bench.py, and synthetic scenes.

The directory name carries a hyphen, so import it with
    importlib.import_module("gaussian-splatterer_b200")
"""
from . import camera, synth  # noqa: F401
from ._lib import EXPORTED_SYMBOLS, LIB_PATH, GsbError, lib  # noqa: F401
from .camera import Camera, CameraSphere, Project  # noqa: F401


def __getattr__(name):
    # torch-dependent modules load lazily so the CPU-only parts import without CUDA
    if name in ("Rasterizer", "Workspaces", "state", "tile_order"):
        from . import rasterizer
        return getattr(rasterizer, name)
    if name in ("Trainer", "Comm", "Window", "ModelSplatsDevice", "ModelSplatsHost", "camera_struct"):
        from . import trainer
        return getattr(trainer, name)
    raise AttributeError(name)
