"""ctypes binding of libb200lz4.so (include/b200lz4.h).

The library is the product; this module only loads it.  There is no Python or CPU codec behind it:
if the shared library is missing or no CUDA device is usable, calls fail loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2LZ4_LIB") or os.path.join(_HERE, "libb200lz4.so")   # B2LZ4_LIB: tuning builds (tools/variants.sh)


class Prefs(C.Structure):
    """b2lz4_prefs — see include/b200lz4.h (mirrors Preferences/FrameInfo of CTypes.hsc:155-232)."""
    _fields_ = [(n, C.c_uint32) for n in (
        "block_size_id", "block_checksum", "content_checksum", "content_size",
        "sub_block_log2", "write_index", "use_index", "block_linked")]


class FrameInfo(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in (
        "block_size_id", "block_independent", "block_checksum", "content_checksum",
        "has_content_size", "has_dict_id", "dict_id", "header_size")] + [
        ("content_size", C.c_uint64), ("block_size", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("n_blocks", C.c_uint32), ("n_chains", C.c_uint32), ("indexed", C.c_uint32),
                ("kernel_launches", C.c_uint32), ("ms_total", C.c_float), ("ms_main", C.c_float),
                ("ms_aux", C.c_float), ("ms_h2d", C.c_float), ("ms_d2h", C.c_float),
                ("bytes_in", C.c_uint64), ("bytes_out", C.c_uint64), ("retries", C.c_uint32),
                ("pad", C.c_uint32)]


# every symbol include/b200lz4.h declares: name -> (restype, argtypes)
_vp, _sz, _u32p, _szp = C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_size_t)
SYMBOLS = {
    "b2lz4_is_error": (C.c_uint, [_sz]),
    "b2lz4_error_name": (C.c_char_p, [_sz]),
    "b2lz4_last_cuda_error": (C.c_char_p, [_vp]),
    "b2lz4_create": (_sz, [C.POINTER(_vp), C.c_int, C.POINTER(Prefs)]),
    "b2lz4_destroy": (_sz, [_vp]),
    "b2lz4_set_prefs": (_sz, [_vp, C.POINTER(Prefs)]),
    "b2lz4_set_devices": (_sz, [_vp, C.POINTER(C.c_int), C.c_int]),
    "b2lz4_set_dict": (_sz, [_vp, _vp, _sz, C.c_uint32]),
    "b2lz4_frame_header_write2": (_sz, [_vp, _sz, C.POINTER(Prefs), C.c_uint64, C.c_uint32]),
    "b2lz4_default_prefs": (None, [C.POINTER(Prefs)]),
    "b2lz4_device_count": (C.c_int, []),
    "b2lz4_compress_bound": (_sz, [_sz, C.POINTER(Prefs)]),
    "b2lz4_decompress_bound": (_sz, [_vp, _sz]),
    "b2lz4_frame_header_write": (_sz, [_vp, _sz, C.POINTER(Prefs), C.c_uint64]),
    "b2lz4_frame_header_parse": (_sz, [_vp, _sz, C.POINTER(FrameInfo)]),
    "b2lz4_compress_device": (_sz, [_vp, _vp, _sz, _vp, _sz]),
    "b2lz4_decompress_device": (_sz, [_vp, _vp, _sz, _vp, _sz, _szp]),
    "b2lz4_compress_host": (_sz, [_vp, _vp, _sz, _vp, _sz]),
    "b2lz4_decompress_host": (_sz, [_vp, _vp, _sz, _vp, _sz, _szp]),
    "b2lz4_pinned_alloc": (_vp, [_sz]),
    "b2lz4_pinned_free": (None, [_vp]),
    "b2lz4_compress_blocks_host": (_sz, [_vp, _vp, _sz, _vp, _sz, C.c_uint]),
    "b2lz4_frame_end": (_sz, [_vp, _vp, _sz]),
    "b2lz4_decompress_blocks_host": (_sz, [_vp, C.POINTER(FrameInfo), _vp, _sz, _vp, _sz, C.c_uint]),
    "b2lz4_content_checksum": (_sz, [_vp, _u32p]),
    "b2lz4_xxh32_device": (_sz, [_vp, _vp, _sz, _u32p]),
    "b2lz4_xxh32_host": (_sz, [_vp, _vp, _sz, _u32p]),
    "b2lz4_last_stats": (_sz, [_vp, C.POINTER(Stats)]),
    # no counterpart in the reference: allocation-free block checksums, csrc/noalloc.cu
    "b2lz4_xxh32_ranges_static": (_sz, [C.c_int, _vp, _vp, _sz, C.POINTER(C.c_uint64), _u32p, C.c_uint32, _u32p]),
    "b2lz4_static_last_error": (C.c_char_p, []),
}

_lib = None


def load():
    """Load libb200lz4.so (built in-tree by __graft_entry__.build() / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `make -C lz4-frame-conduit_b200/csrc` "
                              "(or __graft_entry__.build()); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib
