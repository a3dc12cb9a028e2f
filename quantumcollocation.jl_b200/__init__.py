"""qcb200 -- B200-native evaluator for QuantumCollocation.jl's per-knot collocation hot path.

The product is ``csrc/libqcb200.so`` (hand-written sm_100a CUDA behind the C ABI declared in
``include/qcb200.h``).  This package is the Python host side: ctypes binding, a mirror of the
reference's integrator / objective / constraint objects, synthetic workloads and the
multi-GPU time-slab driver.
"""
from . import isomorphisms, layout, workloads, piccolo, distributed  # noqa: F401
from .layout import KnotLayout, JAC_BLOCK_DIAG, JAC_DENSE_X  # noqa: F401
from .api import Aux, Evaluator, Fidelity, FreePhaseFidelity, Problem  # noqa: F401
from ._lib import QcbError, load as load_library  # noqa: F401
