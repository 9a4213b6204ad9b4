"""cudaneuralrender_b200 -- B200-native (sm_100a) neural-SDF sphere tracer behind the
reference's host interface.  The product is libnsdf.so (csrc/, include/nsdf.h); `host` mirrors
the reference's classes over its C ABI.  Importing `host` needs torch; `_capi` does not."""
from . import _capi  # noqa: F401

__all__ = ["_capi", "host"]


def __getattr__(name):
    if name == "host":
        import importlib
        return importlib.import_module(".host", __name__)
    raise AttributeError(name)
