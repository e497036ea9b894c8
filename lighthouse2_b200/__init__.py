"""RenderCore_B200: a B200-native (sm_100a) wavefront extend/shadow traversal core for Lighthouse 2.

The product is ``lighthouse2_b200/lib/libRenderCore_B200.so`` (C++ CoreAPI plugin + ``extern "C"``
layer over hand-written CUDA, sources in ``lighthouse2_b200/csrc``).  The Python modules here are the
test/bench driver over that C ABI plus procedural scene generators; they never import ``oracle/``.
"""
__version__ = "0.1.0"
