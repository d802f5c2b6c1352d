"""nnstreamer_b200 -- B200-native tensor_transform / tensor_converter hot path (+ tensor_aggregator).

Everything computes on the GPU through libnnstreamer_b200.so (see include/nnstreamer_b200.h);
this package only mirrors the reference element's interface on top of that C-ABI.
"""
from . import _lib  # noqa: F401
from ._lib import NnsbError, lib  # noqa: F401
from .device import Device, Plan  # noqa: F401
from .element import (BatchQueue, Branch, DeviceMemory, DirectVideoDecoder, TensorAggregator,  # noqa: F401
                      TensorConverter, TensorDemux, TensorFilter, TensorMerge, TensorMux,
                      TensorSplit, TensorTransform, device_memory)

__all__ = ["Device", "Plan", "NnsbError", "lib", "TensorTransform", "TensorConverter",
           "TensorAggregator", "DirectVideoDecoder", "TensorMerge", "TensorSplit", "BatchQueue",
           "DeviceMemory", "Branch", "TensorFilter", "TensorMux", "TensorDemux", "device_memory"]
