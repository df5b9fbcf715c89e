"""tetrex_b200 — B200-native (sm_100a) implementation of TetRex's index-probe + verify hot path.

The product is the CUDA library ``tetrex_b200/lib/libtetrex_cuda.so`` behind the C ABI declared in
``include/tetrex_b200.h``; this package is the Python mirror of that ABI used by tests and bench.py.
There is no CPU fallback: importing works anywhere, using it needs the built library and a B200.
"""
from ._abi import LIB_PATH, TetrexError, lib  # noqa: F401
from .api import (HIBF, IBF, dgram_count, hll_sketch_records, Comm, MOL_AA, MOL_NA, RED_BASE, RED_LI, RED_MURPHY, Context, Paths, Regex, Text,  # noqa: F401
                  compute_bitcount, emit_count, peptide_maps)

__version__ = "0.1.0"
