"""ptb200 — Python bindings (ctypes) of the B200-native render path for programmerjake/path-trace.

The product is path-trace_b200/libptb200.so (CUDA kernels + C ABI + C++ host scene API); this
package only binds it for tests and bench.py.
"""
from . import builder, capi, device, scenes  # noqa: F401
