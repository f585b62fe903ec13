"""Kernel LOGIC on the SIMT emulator (tools/emu/): the product's .cu sources compiled with g++ -DB2_EMU, every CUDA thread
a fiber, warp collectives / __syncthreads / inter-warp waits scheduled on the host.  TEST INFRASTRUCTURE ONLY — this is how
the kernels (above all the dataflow decoder, csrc/decode_df.cu, whose warps wait for one another) are exercised in a
container without a GPU; the parity tests proper are the `-m gpu` files, which run the nvcc build on a B200.
The emulated library is built into tests/_build/ and loaded only here; the product never sees it.
Checker: the oracle (oracle/oracle_lz4.c), the committed golden frames, and oracle/_ref when present."""
import ctypes as C
import importlib.util
import os
import random
import subprocess
import sys

import pytest

from conftest import ROOT, gen_bytes, golden_bytes, orc_decode, ref_conduit_decode, _bind_replay, dict_compress, dict_decompress

EMU_LIB = os.path.join(ROOT, "tests", "_build", "libb200lz4_emu.so")
sz = C.c_size_t


def _build_emu():
    subprocess.run(["make", "-C", os.path.join(ROOT, "tools", "emu")], check=True, capture_output=True)
    return EMU_LIB


@pytest.fixture(scope="session")
def emu():
    """A private copy of the host package bound to the emulated library (the real package stays bound to libb200lz4.so)."""
    path = _build_emu()
    pkgdir = os.path.join(ROOT, "lz4-frame-conduit_b200")
    old = os.environ.get("B2LZ4_LIB")
    os.environ["B2LZ4_LIB"] = path
    try:
        spec = importlib.util.spec_from_file_location("b2lz4_emu_pkg", os.path.join(pkgdir, "__init__.py"),
                                                      submodule_search_locations=[pkgdir])
        mod = importlib.util.module_from_spec(spec)
        sys.modules["b2lz4_emu_pkg"] = mod
        spec.loader.exec_module(mod)
        mod.load()
    finally:
        if old is None:
            os.environ.pop("B2LZ4_LIB", None)
        else:
            os.environ["B2LZ4_LIB"] = old
    assert mod.LIB_PATH == path
    return mod


@pytest.fixture(scope="session")
def emu_replay(emu):
    """oracle/conduit_replay.c (the reference conduit's call sequence) linked to the EMULATED LZ4F_* symbols."""
    out = os.path.join(ROOT, "tests", "_build", "libreplay_emu.so")
    src = os.path.join(ROOT, "oracle", "conduit_replay.c")
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-DREPLAY_AGAINST_B200", "-I", os.path.join(ROOT, "include"), "-o", out, src,
                    "-L", os.path.dirname(EMU_LIB), "-lb200lz4_emu", "-Wl,-rpath," + os.path.dirname(EMU_LIB)], check=True)
    return _bind_replay(C.CDLL(out))


def test_emu_decodes_every_golden_frame(emu, manifest):
    """Frames made by the reference library, its conduit call sequence and the stock CLI (tests/golden/make_golden.py):
    linked and independent blocks, every checksum combination — all through the dataflow kernels (no index in any of them)."""
    n = 0
    with emu.Codec(0, use_index=0) as c:
        for name, e in manifest.items():
            if name.startswith("_"):
                continue
            raw = golden_bytes(e["raw"])[:e["size"]]
            for kind, fr in e["frames"].items():
                f = golden_bytes(fr["file"])
                out, used = c.decompress_bytes(f)
                assert out == raw and used == len(f), (name, kind)
                if name in ("datagen200k", "mixed166k") and kind in ("conduit", "cli_B4_BD"):   # linked: dataflow tickets
                    assert c.stats().n_chains > c.stats().n_blocks, (name, kind, "not decoded by pieces")
                n += 1
    assert n >= 35


def test_emu_streaming_conduit_on_goldens(emu, manifest):
    """decompress (Conduit.hsc:598-701) over the emulated LZ4F_* symbols, the input cut into small chunks."""
    for key in ("datagen200k", "mixed166k", "hellos", "empty"):
        e = manifest[key]
        raw = golden_bytes(e["raw"])[:e["size"]]
        for kind in ("conduit", "cli", "cli_B4_BD", "cli_B4_BX"):
            f = golden_bytes(e["frames"][kind]["file"])
            assert b"".join(emu.conduit.decompress(emu.bs_chunks_of(4099, f))) == raw, (key, kind)


def test_emu_roundtrip_shapes_against_oracle(emu, oracle, gen):
    """Encoder kernels on the emulator: every frame decodes with the oracle and with the (emulated) decoder."""
    shapes = [gen_bytes(gen, "loglike", 300000), gen_bytes(gen, "textlike", 200000, 3), bytes(150000),
              gen_bytes(gen, "random", 70000, 5), b"", b"a", b"hello" * 3, gen_bytes(gen, "loglike", 65537)]
    with emu.Codec(0) as c:
        for prefs in (dict(block_size_id=4, sub_block_log2=16, block_checksum=1, content_checksum=1),
                      dict(block_size_id=5, sub_block_log2=18, block_checksum=0, content_checksum=0),
                      dict(block_size_id=7, sub_block_log2=17, write_index=1)):
            c.set_prefs(**prefs)
            for d in shapes:
                fr = c.compress_bytes(d)
                got, used = orc_decode(oracle, fr, len(d) + 16)
                assert got == d, (prefs, len(d))
                back, used2 = c.decompress_bytes(fr)
                assert back == d and used2 == len(fr), (prefs, len(d))


def test_emu_dataflow_vs_oracle_on_corrupt_frames(emu, oracle, gen):
    """Accept / reject must equal the oracle's verdict on damaged blocks (frametest.c:1044-1073 idea), bytes equal when both
    accept.  One documented deviation is common to oracle and product (offset 0 is rejected)."""
    rnd = random.Random(5)
    d = gen_bytes(gen, "loglike", 90000) + bytes(3000) + gen_bytes(gen, "textlike", 50000, 2)
    with emu.Codec(0, use_index=0) as c:
        c.set_prefs(block_size_id=4, sub_block_log2=16, write_index=0)
        good = c.compress_bytes(d)
        seen_bad = seen_ok = 0
        for i in range(60):
            b = bytearray(good)
            for _ in range(rnd.randint(1, 3)):
                b[rnd.randrange(7, len(b))] = rnd.getrandbits(8)
            b = bytes(b)
            want, used = orc_decode(oracle, b, len(d) + 70000)
            try:
                got, used2 = c.decompress_bytes(b, max_out=len(d) + 70000)
            except emu.Lz4FrameError:
                got = None
            if isinstance(want, int):
                assert got is None, (i, want)
                seen_bad += 1
            else:
                assert got == want and used2 == used, i
                seen_ok += 1
        assert seen_bad > 5 and seen_ok > 0


def test_emu_reference_made_linked_frame(emu, emu_replay, ref, gen):
    """A frame the reference conduit produced (linked 64 KB blocks): one-shot decode and the streaming LZ4F_* layer in
    several batches (history carried in front of the batch), under the reference conduit's own decompress loop."""
    if ref is None:
        pytest.skip("oracle/_ref not built")
    d = gen_bytes(gen, "loglike", 9 << 20) + gen_bytes(gen, "random", 100000, 1) + bytes(70000)
    cap = len(d) + len(d) // 100 + 100000
    o = C.create_string_buffer(cap)
    r = ref.ref_conduit_compress(d, len(d), o, cap, 32768, 0)
    frame = o.raw[:r]
    assert frame[:7].hex() == "04224d184040c0"
    with emu.Codec(0, use_index=0) as c:
        got, used = c.decompress_bytes(frame)
        assert got == d and used == len(frame)
    os.environ["B2LZ4_BATCH_MB"] = "4"
    got, used = ref_conduit_decode(emu_replay, frame, len(d) + 64, chunk=50000)       # 3 batches of 4 MiB
    assert got == d and used == len(frame)


def test_emu_linked_block_encoding(emu, oracle, ref, lz4_cli, datagen, gen):
    """blockMode = linked: FLG 0x40 frames whose matches cross block borders.  Every decoder takes them, and
    at 64 KB blocks the size is within the 2 % of the reference's default (linked 64 KB) that independent blocks cannot reach."""
    shapes = [gen_bytes(gen, "loglike", 500000), bytes(200000) + gen_bytes(gen, "random", 70000, 2) + gen_bytes(gen, "textlike", 90000, 4), b"ab" * 7]
    if datagen:
        shapes.append(subprocess.run([datagen, "-g1500000", "-s1"], capture_output=True, check=True).stdout)
    with emu.Codec(0, use_index=0) as c:
        for bid in (4, 5):
            c.set_prefs(block_size_id=bid, sub_block_log2=16 if bid == 4 else 18, block_linked=1, write_index=0, content_checksum=1)
            for d in shapes:
                fr = c.compress_bytes(d)
                assert fr[4] & 0x20 == 0                                  # FLG bit 5 clear: linked (lz4_Frame_format.md:136-138)
                got, used = orc_decode(oracle, fr, len(d) + 16)
                assert got == d and used == len(fr), (bid, len(d))
                assert c.decompress_bytes(fr)[0] == d
                if lz4_cli:
                    assert subprocess.run([lz4_cli, "-d", "-c"], input=fr, capture_output=True).stdout == d
                if ref is not None and len(d) > 400000:
                    cap = len(d) + len(d) // 100 + 100000
                    o = C.create_string_buffer(cap)
                    r = ref.ref_conduit_compress(d, len(d), o, cap, 0, 0)
                    assert len(fr) <= r * 1.02, (bid, len(fr), r)


def test_emu_dictionary_frames_both_ways(emu_replay, ref, lz4_cli, manifest, gen, tmp_path):
    """LZ4F_createCDict / LZ4F_compressBegin_usingCDict / LZ4F_decompress_usingDict (lz4frame.c:478-521, 594-702, 1842-1860):
    reference-made dictionary frames (committed goldens) decode here; frames made here decode with the reference library
    and with `lz4 -d -D`; without the dictionary they must not."""
    dman = manifest["_dict"]
    dic = golden_bytes(dman["file"])
    for name, fr in dman["frames"].items():
        raw = golden_bytes(manifest[name]["raw"])[:manifest[name]["size"]]
        f = golden_bytes(fr["file"])
        assert dict_decompress(emu_replay, f, dic, len(raw) + 64, 7000) == raw, name
        if name in ("datagen200k", "mixed166k"):                                          # same generator as the dictionary:
            assert dict_decompress(emu_replay, f, b"", len(raw) + 64) != raw, name       # the dictionary is really needed
    dfile = tmp_path / "dict"
    dfile.write_bytes(dic)
    for d in (b"", b"x", dic[5000:5300], dic[100:9000] + gen_bytes(gen, "loglike", 100000), gen_bytes(gen, "textlike", 300000, 3)):
        for did in (0, 0xABCD1234):
            ours = dict_compress(emu_replay, d, dic, 4, 0, 1, did)
            assert ours[4] & 0x20 == 0 and bool(ours[4] & 1) == bool(did)
            assert dict_decompress(emu_replay, ours, dic, len(d) + 64, 999) == d
            if ref is not None:
                assert dict_decompress(ref, ours, dic, len(d) + 64, 1000) == d
                indep = dict_compress(ref, d, dic, 4, 1, 1, did)                  # independent blocks + dictionary (what `lz4 -D` writes for one block)
                assert dict_decompress(emu_replay, indep, dic, len(d) + 64, 4096) == d
                theirs = dict_compress(ref, d, dic, 4, 0, 1, did)
                assert ours[:len(theirs[:11]) if did else 7] == theirs[:11 if did else 7]        # same header bytes
                assert dict_decompress(emu_replay, theirs, dic, len(d) + 64) == d
                if 200 < len(d) < 10000:
                    assert len(ours) < len(dict_compress(emu_replay, d, b"", 4, 0, 1, did))   # the dictionary pays
            if lz4_cli:
                r = subprocess.run([lz4_cli, "-d", "-c", "-D", str(dfile)], input=ours, capture_output=True)
                assert r.returncode == 0 and r.stdout == d


def _run_with_env(env, code):
    e = dict(os.environ); e.update(env); e["B2LZ4_LIB"] = EMU_LIB
    p = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-800:] + p.stderr[-1500:]
    return p.stdout


_GOLDEN_LOOP = """
import importlib, json, sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
pkg = importlib.import_module('lz4-frame-conduit_b200')
man = json.load(open('tests/golden/manifest.json'))
with pkg.Codec(0, use_index=0) as c:
    for name in ('datagen200k', 'mixed166k', 'numbers', 'hellos'):
        e = man[name]
        raw = open('tests/golden/' + e['raw'], 'rb').read()[:e['size']]
        for kind, fr in e['frames'].items():
            out, used = c.decompress_bytes(open('tests/golden/' + fr['file'], 'rb').read())
            assert out == raw, (name, kind)
        fr = c.compress_bytes(raw)
        assert c.decompress_bytes(fr)[0] == raw
print('ok')
"""


def test_emu_pooled_engine_contexts_carry_nothing_over(emu_replay, oracle, manifest, gen):
    """Freed LZ4F contexts hand their engine context to a process-wide pool (lz4f_compat.cu: EnginePool); the next LZ4F
    context gets it back with other preferences.  A dictionary, linked-block history, checksum state or block size of the
    stream before must not show in the stream after — in either direction."""
    dic = golden_bytes(manifest["_dict"]["file"])
    a = gen_bytes(gen, "loglike", 400000, 11)
    b = gen_bytes(gen, "textlike", 300000, 12) + bytes(50000)
    cap = 1 << 20
    for rnd in range(2):
        with_dict = dict_compress(emu_replay, a, dic, 4, 0, 1, 77)                   # linked blocks + dictionary + content checksum
        assert dict_decompress(emu_replay, with_dict, dic, len(a) + 64, 5000) == a
        o = C.create_string_buffer(cap)
        r = emu_replay.ref_conduit_compress(b, len(b), o, cap, 16384, 0)             # the plain conduit right after: no dictionary, no checksum
        plain = o.raw[:r]
        got, used = orc_decode(oracle, plain, len(b) + 64)
        assert got == b and used == len(plain), rnd
        got, used = ref_conduit_decode(emu_replay, plain, len(b) + 64, chunk=7001)
        assert got == b and used == len(plain), rnd
        assert dict_decompress(emu_replay, with_dict, dic, len(a) + 64) == a        # and the dictionary stream again after the plain one


def test_emu_dataflow_forced_for_every_frame(emu):
    """B2LZ4_DF=2: independent small-block frames (normally one warp per block) through the dataflow kernels as well."""
    assert "ok" in _run_with_env({"B2LZ4_DF": "2"}, _GOLDEN_LOOP)
    assert "ok" in _run_with_env({"B2LZ4_DF": "0"}, _GOLDEN_LOOP)          # and everything without them


def test_emu_static_scratch_checksums(emu, oracle, gen):
    """b2lz4_xxh32_ranges_static on the emulator (the GPU test is tests/test_zz_no_alloc.py): hashes equal the oracle's, two calls in
    a row share the scratch array, the limit and out-of-buffer ranges are refused with the documented error."""
    lib = emu.load()
    data = gen_bytes(gen, "loglike", 300000) + bytes(5000) + gen_bytes(gen, "random", 60000, 3)
    buf = C.create_string_buffer(data, len(data))
    rnd = random.Random(9)
    for n in (400, 150):
        offs = [rnd.randrange(0, len(data) - 9000) for _ in range(n)]
        lens = [rnd.choice((0, 1, 15, 16, 17, 255, 4096, 8999)) for _ in range(n)]
        a = (C.c_uint64 * n)(*offs); b = (C.c_uint32 * n)(*lens); out = (C.c_uint32 * n)()
        r = lib.b2lz4_xxh32_ranges_static(0, None, C.addressof(buf), len(data), a, b, n, out)
        assert not lib.b2lz4_is_error(r), lib.b2lz4_static_last_error()
        for i in range(n):
            assert out[i] == oracle.orc_xxh32(data[offs[i]:offs[i] + lens[i]], lens[i], 0), (n, i)
    big = 65537
    r = lib.b2lz4_xxh32_ranges_static(0, None, C.addressof(buf), len(data), (C.c_uint64 * big)(), (C.c_uint32 * big)(), big, (C.c_uint32 * big)())
    assert lib.b2lz4_is_error(r) and lib.b2lz4_error_name(r) == b"ERROR_srcSize_tooLarge" and b"B2LZ4_STATIC_MAX_RANGES" in lib.b2lz4_static_last_error()
    r = lib.b2lz4_xxh32_ranges_static(0, None, C.addressof(buf), len(data), (C.c_uint64 * 1)(len(data) - 3), (C.c_uint32 * 1)(4), 1, (C.c_uint32 * 1)())
    assert lib.b2lz4_is_error(r) and b"leaves the buffer" in lib.b2lz4_static_last_error()


_STREAM_LOOP = """
import ctypes as C, os, random, subprocess, sys
lib = C.CDLL(os.environ['B2LZ4_LIB']); sz = C.c_size_t
lib.LZ4F_createDecompressionContext.argtypes = [C.POINTER(C.c_void_p), C.c_uint]; lib.LZ4F_createDecompressionContext.restype = sz
lib.LZ4F_freeDecompressionContext.argtypes = [C.c_void_p]; lib.LZ4F_freeDecompressionContext.restype = sz
lib.LZ4F_decompress.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(sz), C.c_void_p, C.POINTER(sz), C.c_void_p]; lib.LZ4F_decompress.restype = sz
lib.LZ4F_isError.argtypes = [sz]; lib.LZ4F_getErrorName.argtypes = [sz]; lib.LZ4F_getErrorName.restype = C.c_char_p
gen = C.CDLL('tools/libb2gen.so'); gen.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
b = C.create_string_buffer(9 << 20); gen.b2gen_loglike(b, 9 << 20, 21)
d = b.raw + os.urandom(300000) + bytes(1 << 20)
rnd = random.Random(4)
for flags in (['-B4'], ['-B4', '-BD'], ['-B6', '-BX'], ['-B7', '--no-frame-crc']):
    fr = subprocess.run(['oracle/_ref/lz4', '-c', *flags], input=d, capture_output=True, check=True).stdout
    for srcmax, dstmax in ((65536, 65536), (1 << 20, 9000), (3000, 2 << 20)):
        ctx = C.c_void_p(); assert lib.LZ4F_createDecompressionContext(C.byref(ctx), 100) == 0
        out = bytearray(); pos = 0; dst = C.create_string_buffer(dstmax); src = C.create_string_buffer(fr, len(fr))
        while True:
            take = min(len(fr) - pos, rnd.randint(1, srcmax)); s = sz(take); o = sz(rnd.randint(0 if take else 1, dstmax))
            h = lib.LZ4F_decompress(ctx, dst, C.byref(o), C.addressof(src) + pos, C.byref(s), None)
            assert not lib.LZ4F_isError(h), lib.LZ4F_getErrorName(h)
            out += dst.raw[:o.value]; pos += s.value
            if h == 0: break
        lib.LZ4F_freeDecompressionContext(ctx)
        assert bytes(out) == d and pos == len(fr), (flags, srcmax, dstmax, len(out), pos)
print('ok')
"""


def test_emu_streaming_decode_several_batches_random_slices(emu, lz4_cli):
    """LZ4F_decompress (lz4frame.c:1380-1840 contract) over stock-CLI frames that take several GPU batches (B2LZ4_BATCH_MB=4), fed
    and drained in random slices: the layer stages and launches the next batch while the caller still drains the one before."""
    if lz4_cli is None:
        pytest.skip("oracle/_ref/lz4 not present")
    assert "ok" in _run_with_env({"B2LZ4_BATCH_MB": "4"}, _STREAM_LOOP)


def test_emu_lane_order_does_not_matter(emu):
    """The lanes of a warp run in reverse / shuffled order between collectives: a missing __syncwarp() shows as a wrong byte."""
    for order in ("reverse", "random"):
        assert "ok" in _run_with_env({"B2_EMU_LANE_ORDER": order}, _GOLDEN_LOOP)


def test_emu_no_deadlock_with_one_resident_cta(emu):
    """Dataflow forward progress must not depend on how many CTAs are resident: one CTA at a time, CTAs dispatched in
    order as earlier ones retire (the emulator aborts with 'no progress' on a wait nobody can satisfy)."""
    assert "ok" in _run_with_env({"B2_EMU_SMS": "1", "B2_EMU_RESIDENT_THREADS": "256"}, _GOLDEN_LOOP)
