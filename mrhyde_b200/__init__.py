"""mrhyde_b200 — B200-native finite-element assembly engine (MrHyDE hot path).

The product is `libmrhyde_b200.so` (C++ host + hand-written sm_100a CUDA behind the C ABI of
include/mrhyde_b200.h).  This package only builds it (`build.py`) and binds it with ctypes (`engine.py`) for the
tests and bench.py.  There is no CPU fallback: without the CUDA library or without a GPU every assemble call fails.
"""
from .engine import Engine, MrhError, lib_path, load_library  # noqa: F401
