import ctypes as C
import importlib
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# the LZ4F_* layer batches this many MiB per GPU call; small here so that modest test inputs span several batches
os.environ.setdefault("B2LZ4_BATCH_MB", "8")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


sz = C.c_size_t


@pytest.fixture(scope="session")
def oracle():
    """oracle/liboracle_lz4.so — the CPU restatement (test infrastructure only)."""
    path = os.path.join(ROOT, "oracle", "liboracle_lz4.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, capture_output=True)
    lib = C.CDLL(path)
    lib.orc_xxh32.restype = C.c_uint32
    lib.orc_xxh32.argtypes = [C.c_char_p, sz, C.c_uint32]
    lib.orc_block_compress.restype = C.c_int
    lib.orc_block_compress.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
    lib.orc_block_decompress.restype = C.c_int
    lib.orc_block_decompress.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, sz]
    lib.orc_frame_decompress.restype = sz
    lib.orc_frame_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.POINTER(sz)]
    lib.orc_frame_compress.restype = sz
    lib.orc_frame_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_void_p]
    lib.orc_frame_header_write.restype = sz
    lib.orc_frame_header_write.argtypes = [C.c_char_p, sz, C.c_void_p]
    lib.orc_frame_header_parse.restype = sz
    lib.orc_frame_header_parse.argtypes = [C.c_char_p, sz, C.c_void_p]
    return lib


class OrcDesc(C.Structure):
    _fields_ = [("block_size_id", C.c_int), ("block_independent", C.c_int), ("content_checksum", C.c_int),
                ("block_checksum", C.c_int), ("has_content_size", C.c_int), ("content_size", C.c_uint64),
                ("has_dict_id", C.c_int), ("dict_id", C.c_uint32)]


def orc_decode(lib, frame, cap):
    """Decode one frame with the oracle; returns (bytes | error-code int, consumed)."""
    out = C.create_string_buffer(max(cap, 1))
    consumed = sz(0)
    r = lib.orc_frame_decompress(frame, len(frame), out, cap, C.byref(consumed))
    if r > (1 << 63):
        return -((1 << 64) - r), 0          # negative error number
    return out.raw[:r], consumed.value


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref/libref.so — the UNMODIFIED reference library + our conduit replay (built by
    `make -C oracle ref`; travels to the GPU box as a built file).  None when absent."""
    path = os.path.join(ROOT, "oracle", "_ref", "libref.so")
    if not os.path.exists(path):
        if os.path.isdir("/root/reference/lz4/lib"):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, capture_output=True)
        else:
            return None
    lib = C.CDLL(path)
    for name in ("ref_conduit_compress", "ref_conduit_compress_yi"):
        getattr(lib, name).restype = sz
    lib.ref_conduit_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz, sz]
    lib.ref_conduit_compress_yi.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz]
    lib.ref_conduit_decompress.restype = sz
    lib.ref_conduit_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz, C.POINTER(sz)]
    lib.ref_frame_decompress.restype = sz
    lib.ref_frame_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.POINTER(sz)]
    lib.ref_frame_compress.restype = sz
    lib.ref_frame_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz] + [C.c_int] * 6 + [sz]
    lib.ref_xxh32.restype = C.c_uint32
    lib.ref_xxh32.argtypes = [C.c_char_p, sz, C.c_uint32]
    lib.ref_xxh32_stream.restype = C.c_uint32
    lib.ref_xxh32_stream.argtypes = [C.c_char_p, sz, C.c_uint32, sz]
    lib.LZ4F_isError.restype = C.c_uint
    lib.LZ4F_isError.argtypes = [sz]
    lib.LZ4F_getErrorName.restype = C.c_char_p
    lib.LZ4F_getErrorName.argtypes = [sz]
    lib.LZ4_compress_default.restype = C.c_int
    lib.LZ4_compress_default.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int]
    lib.LZ4_decompress_safe.restype = C.c_int
    lib.LZ4_decompress_safe.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int]
    if hasattr(lib, "ref_frame_compress_dict"):
        lib.ref_frame_compress_dict.restype = sz
        lib.ref_frame_compress_dict.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_char_p, sz, C.c_int, C.c_int, C.c_int, C.c_uint, sz]
        lib.ref_frame_decompress_dict.restype = sz
        lib.ref_frame_decompress_dict.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_char_p, sz, sz]
    return lib


def _bind_replay(lib):
    lib.ref_conduit_compress.restype = sz
    lib.ref_conduit_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz, sz]
    lib.ref_conduit_compress_yi.restype = sz
    lib.ref_conduit_compress_yi.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz]
    lib.ref_conduit_decompress.restype = sz
    lib.ref_conduit_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz, C.POINTER(sz)]
    lib.ref_frame_decompress.restype = sz
    lib.ref_frame_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.POINTER(sz)]
    lib.ref_frame_compress.restype = sz
    lib.ref_frame_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz] + [C.c_int] * 6 + [sz]
    lib.ref_sizeof_frameinfo.restype = sz
    lib.ref_sizeof_preferences.restype = sz
    lib.ref_frame_compress_dict.restype = sz
    lib.ref_frame_compress_dict.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_char_p, sz, C.c_int, C.c_int, C.c_int, C.c_uint, sz]
    lib.ref_frame_decompress_dict.restype = sz
    lib.ref_frame_decompress_dict.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_char_p, sz, sz]
    return lib


def dict_compress(lib, data, dic, bsid=4, indep=0, cc=1, dict_id=0, chunk=16384):
    """LZ4F_createCDict + LZ4F_compressBegin_usingCDict + updates + end (oracle/conduit_replay.c) on `lib`'s LZ4F_* symbols."""
    cap = len(data) + len(data) // 50 + 70000
    out = C.create_string_buffer(cap)
    r = lib.ref_frame_compress_dict(data, len(data), out, cap, dic, len(dic), bsid, indep, cc, dict_id, chunk)
    assert r < (1 << 62), f"dict compress failed with code {(1 << 64) - r}"
    return out.raw[:r]


def dict_decompress(lib, frame, dic, cap, chunk=0):
    """LZ4F_decompress_usingDict in a loop; returns the bytes or a negative error number."""
    out = C.create_string_buffer(cap + 1)
    r = lib.ref_frame_decompress_dict(frame, len(frame), out, cap, dic, len(dic), chunk)
    if r > (1 << 62):
        return -((1 << 64) - r)
    return out.raw[:r]


def build_replay_b200():
    """tests/_build/libreplay_b200.so: oracle/conduit_replay.c (the C replay of Conduit.hsc's call sequence)
    compiled against include/lz4frame.h and linked to the PRODUCT library — the reference's conduit driving
    our LZ4F_* layer.  Returns the path."""
    out_dir = os.path.join(ROOT, "tests", "_build")
    out = os.path.join(out_dir, "libreplay_b200.so")
    srcs = [os.path.join(ROOT, "oracle", "conduit_replay.c"), os.path.join(ROOT, "include", "lz4frame.h")]
    libdir = os.path.join(ROOT, "lz4-frame-conduit_b200")
    if not os.path.exists(out) or any(os.path.getmtime(p) > os.path.getmtime(out) for p in srcs):
        os.makedirs(out_dir, exist_ok=True)
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-DREPLAY_AGAINST_B200", "-I", os.path.join(ROOT, "include"),
                        "-o", out, srcs[0], "-L", libdir, "-lb200lz4", "-Wl,-rpath," + libdir], check=True)
    return out


@pytest.fixture(scope="session")
def replay(pkg):
    """The conduit replay bound to libb200lz4.so's LZ4F_* symbols."""
    pkg.load()
    return _bind_replay(C.CDLL(build_replay_b200()))


def ref_conduit_decode(lib, frame, cap, chunk=0):
    out = C.create_string_buffer(max(cap, 1))
    consumed = sz(0)
    r = lib.ref_conduit_decompress(frame, len(frame), out, cap, chunk, C.byref(consumed))
    if r > (1 << 63):
        return -((1 << 64) - r), 0
    return out.raw[:r], consumed.value


@pytest.fixture(scope="session")
def lz4_cli():
    """The stock lz4 CLI built from the reference sources (oracle/_ref/lz4), or None."""
    path = os.path.join(ROOT, "oracle", "_ref", "lz4")
    return path if os.path.exists(path) else None


@pytest.fixture(scope="session")
def datagen():
    path = os.path.join(ROOT, "oracle", "_ref", "datagen")
    return path if os.path.exists(path) else None


@pytest.fixture(scope="session")
def manifest():
    with open(os.path.join(GOLDEN, "manifest.json")) as f:
        return json.load(f)


def golden_bytes(name):
    with open(os.path.join(GOLDEN, name), "rb") as f:
        return f.read()


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module("lz4-frame-conduit_b200")


@pytest.fixture(scope="session")
def gen():
    """tools/libb2gen.so — synthetic input generators (bench/test utility)."""
    path = os.path.join(ROOT, "tools", "libb2gen.so")
    if not os.path.exists(path):
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", path, os.path.join(ROOT, "tools", "b2gen.c")], check=True)
    lib = C.CDLL(path)
    for n in ("b2gen_loglike", "b2gen_random", "b2gen_textlike"):
        getattr(lib, n).argtypes = [C.c_void_p, sz, C.c_uint64]
        getattr(lib, n).restype = None
    return lib


def gen_bytes(lib, kind, n, seed=0x10C5):
    buf = C.create_string_buffer(n) if n else C.create_string_buffer(1)
    getattr(lib, "b2gen_" + kind)(buf, n, seed)
    return buf.raw[:n]
