"""B200-native prompted-CLIP training step (PromptSRC / CoOp / CoCoOp / MaPLe hot path).

Package layout (only what the path needs):
  csrc/     hand-written sm_100a kernels + the C ABI (include/psrc_b200.h)
  build.py  in-tree nvcc build of lib/libpsrc_b200.so
  cabi.py   ctypes binding of the C ABI          ops.py   tensor-level wrappers
The importable alias is ``psrc_b200`` (repo root shim), because this directory name
is not a Python identifier.
"""
from . import cabi  # noqa: F401
from . import ops  # noqa: F401
from . import synthetic  # noqa: F401

__all__ = ["cabi", "ops"]
