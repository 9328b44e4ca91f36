"""hgpumap -- host-side mirror of the reference's K-means map path for B200.

Python here is plumbing only: ``capi`` binds the C ABI of ``libhgpumap.so``
(include/hgpumap.h) with ctypes, ``wire`` speaks the Hadoop Pipes binary
protocol (to drive Pipes children without a JVM), ``synth`` restates the
benchmark's synthetic workload in numpy.  There is no CPU implementation of
the map path in this package: without the CUDA library every compute call
raises.
"""
from . import wire  # noqa: F401
from .capi import (  # noqa: F401
    FMT_F32BE,
    FMT_F32LE,
    FMT_TEXT,
    KERNEL_AUTO,
    KERNEL_EXACT,
    KERNEL_GRID2D,
    KERNEL_TENSOR,
    GpuMapSlot,
    HgmError,
    library_path,
    load_library,
)
