"""gelex_b200 — B200-native GRM engine behind the `gelex grm` interface.

Only what the hot path needs lives here: `csrc/` (CUDA kernels + C ABI + C++ host engine/CLI) and thin ctypes
marshalling (`_lib.py`, `engine.py`).  Importing this package never imports anything under oracle/.
"""
from . import _lib  # noqa: F401
from ._lib import (Accumulator, GrmError, KIND_ADD, KIND_DOM, GROUP_WHOLE, PREC_BF16, PREC_FP16, PREC_FP64,  # noqa: F401
                   METHODS, decode_chunk, standardise_chunk, assoc_chunk, genotype_chunk, synth_bed_host, synth_bed_device, device_count, tri, ipc_open, ipc_close)
