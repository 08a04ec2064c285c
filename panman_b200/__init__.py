"""panman_b200 -- B200-native Fitch / Sankoff small parsimony for PanMAN's MSA construction path.

The product is the C-ABI shared library ``libpmfs.so`` (include/pmfs.h, sources in panman_b200/csrc/).  This package is
the thin Python mirror used by the tests and bench.py: ctypes bindings, nibble packing helpers and the column-shard
helpers for multi-GPU runs.  There is no CPU implementation here; importing :mod:`panman_b200.api` without the built
library raises.
"""
from .api import (  # noqa: F401
    FITCH, SANKOFF, MUT_DTYPE, NUCMUT_DTYPE, Context, PmfsError, SynthSpec, build_library, library_path,
    pack_nibbles, unpack_nibbles,
)
