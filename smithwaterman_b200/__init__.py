"""swb200 -- B200-native Smith-Waterman / Needleman-Wunsch engine (hot path of yamule/smithwaterman).

Host mirror of the reference API in `engine`; CUDA kernels + C ABI in `csrc/` (include/swb200.h).
"""
from .engine import (ALIGN_GLOBAL, ALIGN_GLOCAL, ALIGN_LOCAL, Batch, PositionSpecificMatrix, SeqData, SequenceAlignment,  # noqa: F401
                     SubstitutionMatrix, SwbError, cli_record, dpx_peak, pack_sequences, rust_f32_display)
from ._lib import build, lib  # noqa: F401
from . import search  # noqa: F401,E402
