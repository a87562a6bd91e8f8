"""renderengine_b200 — B200-native (sm_100a CUDA) raster hot path behind RenderEngine's
Scene / Camera / MaterialSystem / ShadowSystem API surface.

The product is the C-ABI shared library built from renderengine_b200/csrc (include/reb.h) plus the
C++ host facade in renderengine_b200/host.  This Python package is the thin ctypes mirror used by
tests and bench.py; it has NO CPU fallback: every render call goes through libreb.so and raises if
the library or a CUDA device is missing.
"""
__version__ = "0.1.0"
