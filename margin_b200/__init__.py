"""margin_b200 — B200-native pair-HMM hot path of MarginPolish (UCSC-nanopore-cgl/margin).

Only what the path needs: ``csrc/`` (CUDA kernels + the C ABI of include/margin_b200.h),
``abi`` (ctypes mirror), ``model`` (params JSON -> flattened state machines), ``symbols``
(RLE / Symbol strings / the poa.c caller contract), ``aligner`` (host mirror of
inc/pairwiseAligner.h on top of the C ABI) and ``synth`` (BASELINE.json's synthetic workloads).
"""
from . import abi, model, symbols  # noqa: F401

__version__ = "0.1.0"
