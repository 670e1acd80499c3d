"""taichi_b200 -- B200-native (sm_100a) replacement for Taichi's sparse-SNode struct-for /
MLS-MPM hot path.  Everything computes in libb200mpm.so (hand-written CUDA behind a C ABI,
include/b200mpm.h); this package is the thin host-side mirror of the reference's Python
surface for that path.  There is no CPU or PyTorch fallback: importing fails loudly if the
shared library is missing."""
from . import _lib  # noqa: F401  (raises ImportError if the CUDA library is not built)
from .mpm import Runtime, NdArray, AotModule, Kernel, Mpm3D  # noqa: F401
from .lang import *  # noqa: F401,F403
from . import lang as _lang

del root  # noqa: F821  (ti.root is looked up live: it only exists after ti.init)


def __getattr__(name):
    if name == "root":
        return _lang.root
    raise AttributeError(name)
