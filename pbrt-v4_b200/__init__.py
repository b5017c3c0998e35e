"""pbrt-v4_b200: B200-native wavefront path-tracing backend for pbrt-v4 (ISET fork).

The product is libb200pt.so (hand-written sm_100a CUDA behind the C ABI of
include/b200pt.h).  This package is the thin Python host side: a ctypes binding
(capi), the host mirror of the reference's WavefrontAggregate interface
(aggregate.B200Aggregate) and procedural scene / ray-batch generators (scenes).
There is no CPU path: everything that computes raises if the library or a GPU
is missing.
"""
from . import capi, scenes, render, parallel  # noqa: F401
from .aggregate import B200Aggregate  # noqa: F401
from .render import B200Renderer, RenderScene  # noqa: F401
from ._build import build  # noqa: F401

__all__ = ["capi", "scenes", "B200Aggregate", "build"]
