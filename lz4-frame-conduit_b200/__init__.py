"""lz4-frame-conduit_b200 — B200-native LZ4 frame codec behind the Codec.Compression.LZ4.Conduit API.

Import with importlib (the directory name carries the reference's hyphens):
    pkg = importlib.import_module("lz4-frame-conduit_b200")
"""
from ._lib import load, LIB_PATH, SYMBOLS, Prefs, FrameInfo, Stats          # noqa: F401
from .codec import (Codec, Lz4FrameError, compress, decompress, bs_chunks_of,   # noqa: F401
                    default_prefs, frame_header_parse, frame_header_write, compress_bound)
from . import conduit                                                        # noqa: F401  (streaming mirror over the LZ4F_* symbols)
