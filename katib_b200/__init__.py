"""katib_b200 -- B200-native arithmetic for Katib's goptuna suggestion service, behind a C ABI.

The product is katib_b200/libkatib_b200.so (hand-written sm_100a CUDA, see include/katib_b200.h).  This
package is only the ctypes view of it used by the tests, bench.py and the host-side service mirror;
importing it never falls back to a CPU implementation: if the shared library is missing, loading fails.
"""
from ._lib import Context, KbError, Service, build, header_symbols, lib, lib_path, tpe_scores_noalloc  # noqa: F401

KB_UNIFORM, KB_DISCRETE, KB_INT, KB_STEP_INT, KB_CATEGORICAL = range(5)
