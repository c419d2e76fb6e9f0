"""rvv-lut-vq_b200: B200-native (sm_100a) LUT vector-quantization hot path behind the reference's operator API.

Layout: ``csrc/`` (hand-written CUDA kernels + the C-ABI of include/lutvq.h, built into liblutvq.so),
``ops.py`` / ``linear.py`` / ``utils.py`` (host-side mirror of the reference's Python interface),
``sharded.py`` (output-channel-sharded decode over torch.distributed), ``synth.py`` (synthetic layers)."""
from . import _lib
from ._lib import LutvqError, build, lib
from .utils import get_int_dtype, pack_int_data, unpack_int_data

__all__ = ["_lib", "LutvqError", "build", "lib", "get_int_dtype", "pack_int_data", "unpack_int_data"]


def __getattr__(name):
    # torch-facing modules are imported lazily so that ABI-only users (ctypes) do not pay for them
    if name in ("ops", "linear", "synth", "sharded", "decode", "hf_loader", "traffic", "kernel_selector", "torch_ops", "pq4"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    if name == "QuantizedLinear":
        from .linear import QuantizedLinear
        return QuantizedLinear
    raise AttributeError(name)
