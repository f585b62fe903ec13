"""CPU tests of the host side: the C-ABI library loads and exports every declared symbol, the
host-only frame helpers agree with the oracle and the golden headers, error naming follows the
reference's convention.  No compute call is made here (there is no GPU in this container)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, OrcDesc, golden_bytes


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load()
    hdr = open(os.path.join(ROOT, "include", "b200lz4.h")).read()
    declared = set(re.findall(r"\b(b2lz4_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "header parse failed"
    assert declared == set(pkg.SYMBOLS), declared ^ set(pkg.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_error_convention(pkg):
    lib = pkg.load()
    # lz4frame.c:249-252: error iff code > (size_t)-maxCode ; names as in lz4frame.h:503-524
    assert not lib.b2lz4_is_error(0) and not lib.b2lz4_is_error(1 << 40)
    code = (1 << 64) - 16
    assert lib.b2lz4_is_error(code) and lib.b2lz4_error_name(code) == b"ERROR_decompressionFailed"
    assert lib.b2lz4_error_name((1 << 64) - 11) == b"ERROR_dstMaxSize_tooSmall"
    assert lib.b2lz4_error_name((1 << 64) - 18) == b"ERROR_contentChecksum_invalid"
    assert lib.b2lz4_error_name((1 << 64) - 20) == b"ERROR_cudaFailure"


def test_error_names_match_reference_library(pkg, ref):
    if ref is None:
        pytest.skip("oracle/_ref not built")
    lib = pkg.load()
    for e in range(1, 20):
        code = (1 << 64) - e
        assert lib.b2lz4_error_name(code) == ref.LZ4F_getErrorName(code), e


def test_bs_chunks_of(pkg):
    # the only literal known-answer of the reference package (test/Main.hs:56-58)
    assert pkg.bs_chunks_of(3, b"abc123def4567") == [b"abc", b"123", b"def", b"456", b"7"]
    assert pkg.bs_chunks_of(5, b"") == [b""]            # `BS.length bs <= chunkSize = [bs]` comes first (Conduit.hsc:428-433)
    assert pkg.bs_chunks_of(5, b"abc") == [b"abc"] and pkg.bs_chunks_of(3, b"abc") == [b"abc"]
    with pytest.raises(ValueError):
        pkg.bs_chunks_of(0, b"x")


def test_header_write_matches_reference_goldens(pkg, manifest):
    p = pkg.default_prefs
    # SURVEY.md appendix A golden header bytes (vendored lib v1.9.2)
    assert pkg.frame_header_write(p(block_size_id=7)).hex() == "04224d18607073"            # lz4 -c --no-frame-crc
    assert pkg.frame_header_write(p(block_size_id=7, content_checksum=1)).hex() == "04224d186470b9"   # lz4 -c
    assert pkg.frame_header_write(p(block_size_id=4, content_checksum=1, block_checksum=1)).hex() == \
        manifest["numbers"]["frames"]["cli_B4_BX"]["head"]
    # contentSize variant: compare with the reference-made frame header
    fr = golden_bytes(manifest["numbers"]["frames"]["lib_B5_cs"]["file"])
    n = manifest["numbers"]["size"]
    assert pkg.frame_header_write(p(block_size_id=5, content_checksum=1, content_size=1), n) == fr[:15]


def test_header_write_matches_oracle(pkg, oracle):
    for bsid in (4, 5, 6, 7):
        for cc in (0, 1):
            for bc in (0, 1):
                for cs in (0, 12345678901):
                    d = OrcDesc(bsid, 1, cc, bc, 1 if cs else 0, cs, 0, 0)
                    out = C.create_string_buffer(19)
                    r = oracle.orc_frame_header_write(out, 19, C.byref(d))
                    ours = pkg.frame_header_write(pkg.default_prefs(block_size_id=bsid, content_checksum=cc,
                                                                    block_checksum=bc, content_size=1 if cs else 0), cs)
                    assert ours == out.raw[:r]


def test_header_parse_goldens_and_errors(pkg, manifest):
    for name, e in manifest.items():
        if name.startswith("_"):
            continue
        for kind, fr in e["frames"].items():
            b = golden_bytes(fr["file"])
            info, hs = pkg.frame_header_parse(b)
            assert hs in (7, 15)
            if kind == "conduit":
                assert (info.block_independent, info.block_size, info.content_checksum, info.block_checksum) == (0, 65536, 0, 0)
            if kind == "lib_B5_cs" and e["size"]:
                assert info.has_content_size == 1 and info.content_size == e["size"] and info.block_size == 262144
    good = bytes.fromhex("04224d184040c0")
    E = pkg.Lz4FrameError
    for bad, name in [(good[:6], "ERROR_frameHeader_incomplete"),
                      (b"\x05" + good[1:], "ERROR_frameType_unknown"),
                      (good[:6] + b"\xc1", "ERROR_headerChecksum_invalid"),
                      (bytes.fromhex("04224d188040c0"), "ERROR_headerVersion_wrong"),
                      (bytes.fromhex("04224d184240c0"), "ERROR_reservedFlag_set"),
                      (bytes.fromhex("04224d184030c0"), "ERROR_maxBlockSize_invalid"),
                      (bytes.fromhex("04224d184041c0"), "ERROR_reservedFlag_set")]:
        with pytest.raises(E) as ei:
            pkg.frame_header_parse(bad)
        assert ei.value.name == name, (bad.hex(), ei.value.name)


def test_compress_bound_covers_stored_worst_case(pkg):
    for n in (0, 1, 65536, (4 << 20) + 1, 1 << 30):
        for bsid in (4, 7):
            b = pkg.compress_bound(n, pkg.default_prefs(block_size_id=bsid, block_checksum=1, content_checksum=1))
            bs = 65536 << (2 * (bsid - 4))
            nb = (n + bs - 1) // bs
            assert b >= 19 + n + nb * 8 + 8


def test_no_cpu_fallback_without_gpu(pkg):
    """The product path must fail loudly when no device is usable (no CPU codec behind the API)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.Lz4FrameError) as ei:
        pkg.Codec()
    assert "cudaFailure" in ei.value.name


def test_product_does_not_reference_oracle():
    """Nothing under the package may import, link or execute oracle/."""
    pdir = os.path.join(ROOT, "lz4-frame-conduit_b200")
    for dp, _, files in os.walk(pdir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in txt and "oracle_lz4" not in txt and "orc_" not in txt, os.path.join(dp, f)


# ------------------------------------------------------------------ Tier 1: the LZ4F_* symbols Conduit.hsc binds
LZ4F_SYMBOLS = ["LZ4F_isError", "LZ4F_getErrorName", "LZ4F_createCompressionContext", "LZ4F_freeCompressionContext",
                "LZ4F_compressBegin", "LZ4F_compressBound", "LZ4F_compressUpdate", "LZ4F_compressEnd",
                "LZ4F_createDecompressionContext", "LZ4F_freeDecompressionContext", "LZ4F_getFrameInfo",
                "LZ4F_decompress"]          # Conduit.hsc:150-151,199,181,292,302,311,321,568,545,579,591


def test_lz4f_symbols_exported_and_declared(pkg):
    lib = pkg.load()
    hdr = open(os.path.join(ROOT, "include", "lz4frame.h")).read()
    declared = set(re.findall(r"\b(LZ4F_[A-Za-z]+)\s*\(", hdr))
    assert set(LZ4F_SYMBOLS) <= declared
    for name in declared:
        assert hasattr(lib, name), name


def test_lz4f_struct_layouts_match_reference(replay, ref):
    # CTypes.hsc:224-232 pokes these structs field by field: sizes 32 / 56
    assert replay.ref_sizeof_frameinfo() == 32 and replay.ref_sizeof_preferences() == 56
    if ref is not None:
        ref.ref_sizeof_frameinfo.restype = ref.ref_sizeof_preferences.restype = C.c_size_t
        assert ref.ref_sizeof_frameinfo() == 32 and ref.ref_sizeof_preferences() == 56


def test_lz4f_contexts_and_header_parse_without_gpu(pkg, manifest):
    """Context life cycle, LZ4F_compressBound and LZ4F_getFrameInfo are host-only; compute entry points must fail
    loudly (ERROR_cudaFailure) when no device is usable — there is no CPU codec behind them."""
    lib = pkg.load()
    sz = C.c_size_t
    lib.LZ4F_createDecompressionContext.restype = sz
    lib.LZ4F_createDecompressionContext.argtypes = [C.POINTER(C.c_void_p), C.c_uint]
    lib.LZ4F_freeDecompressionContext.restype = sz
    lib.LZ4F_freeDecompressionContext.argtypes = [C.c_void_p]
    lib.LZ4F_getFrameInfo.restype = sz
    lib.LZ4F_getFrameInfo.argtypes = [C.c_void_p, C.c_void_p, C.c_char_p, C.POINTER(sz)]
    lib.LZ4F_compressBound.restype = sz
    lib.LZ4F_compressBound.argtypes = [sz, C.c_void_p]
    lib.LZ4F_getErrorName.restype = C.c_char_p
    lib.LZ4F_getErrorName.argtypes = [sz]
    assert lib.LZ4F_freeDecompressionContext(None) == 0                       # lz4frame.c:1081
    d = C.c_void_p()
    assert lib.LZ4F_createDecompressionContext(C.byref(d), 100) == 0
    info = (C.c_uint32 * 8)()
    frame = golden_bytes(manifest["hellos"]["frames"]["cli"]["file"])
    n = sz(7)
    r = lib.LZ4F_getFrameInfo(d, info, frame, C.byref(n))
    assert r == 4 and n.value == 7
    want, _ = pkg.frame_header_parse(frame)
    assert info[0] == want.block_size_id and info[1] == 1 and info[2] == 1    # CLI default: independent, content checksum (lz4io.c:159-175)
    n = sz(7)
    assert lib.LZ4F_getFrameInfo(d, info, frame, C.byref(n)) == 4 and n.value == 0   # already decoded: nothing consumed
    lib.LZ4F_freeDecompressionContext(d)
    d = C.c_void_p()
    lib.LZ4F_createDecompressionContext(C.byref(d), 100)
    n = sz(6)
    assert lib.LZ4F_getErrorName(lib.LZ4F_getFrameInfo(d, info, frame, C.byref(n))) == b"ERROR_frameHeader_incomplete"
    n = sz(7)
    assert lib.LZ4F_getErrorName(lib.LZ4F_getFrameInfo(d, info, b"\x00" * 7, C.byref(n))) == b"ERROR_frameType_unknown"
    lib.LZ4F_freeDecompressionContext(d)
    # bound: a batch of buffered input + the call's bytes, all stored, + EndMark + checksum
    b = lib.LZ4F_compressBound(19 + 16384, None)
    assert b >= (8 << 20) - 1 + 19 + 16384 + 8 and not pkg.load().b2lz4_is_error(b)


def test_lz4f_compress_fails_loudly_without_gpu(replay, pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    out = C.create_string_buffer(64 << 20)
    r = replay.ref_conduit_compress(b"hello" * 100, 500, out, len(out), 0, 0)
    assert pkg.load().b2lz4_error_name(r) == b"ERROR_cudaFailure"


def test_cli_builds_and_fails_loudly_without_gpu(pkg):
    """b200lz4c (the haskell-lz4c twin) has no CPU codec either: without a device it must exit non-zero with the
    reference's error text shape (Conduit.hsc:145-160)."""
    import subprocess
    from conftest import _have_gpu
    cli = os.path.join(ROOT, "lz4-frame-conduit_b200", "b200lz4c")
    if not os.path.exists(cli):
        subprocess.run(["make", "-C", os.path.join(ROOT, "lz4-frame-conduit_b200", "csrc")], check=True, capture_output=True)
    assert b"Compress or decompress .lz4 files" in subprocess.run([cli, "--help"], capture_output=True).stdout   # app/Main.hs:40
    if _have_gpu():
        return
    p = subprocess.run([cli], input=b"hello", capture_output=True)
    assert p.returncode != 0 and p.stdout == b"" and b"lz4frame error: ERROR_cudaFailure" in p.stderr
