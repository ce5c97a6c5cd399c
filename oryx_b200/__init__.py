"""oryx_b200 -- B200-native (sm_100a) ALS serving/speed hot path behind a C ABI.

The product is ``liboryx_b200.so`` (hand-written CUDA kernels + the ``extern "C"`` surface declared
in ``include/oryx_b200.h``).  This Python package is only a ctypes binding over that ABI for tests,
the benchmark and Python callers; it contains no compute and has no CPU fallback: importing
``oryx_b200.native`` without the built library raises.
"""
from .native import (COSINE, DOT, X, Y, Exchange, Model, OryxError, PeerStores, SingularMatrixSolverException, Solver,
                     build, lib, lsh_plan, merge_topn_batch_device, merge_topn_device, merge_topn_packed, packed_record_bytes)

__all__ = ["COSINE", "DOT", "X", "Y", "Exchange", "Model", "OryxError", "PeerStores", "SingularMatrixSolverException",
           "Solver", "build", "lib", "lsh_plan", "merge_topn_batch_device", "merge_topn_device", "merge_topn_packed", "packed_record_bytes"]
