"""GPU tests of the LZ4F_* layer (include/lz4frame.h) — the symbols the UNMODIFIED Conduit.hsc binds.

The driver is oracle/conduit_replay.c, the C replay of the Haskell conduits' exact call sequences
(Conduit.hsc:457-533 compress, :364-425 compressYieldImmediately, :598-701 decompress), compiled a second time
against our header and linked to libb200lz4.so (conftest.build_replay_b200).  The same replay linked to the
unmodified reference library (oracle/_ref/libref.so) and the oracle are the checkers.  These restate the
reference's own test groups (test/Main.hs:60-119): compress -> reference decodes; reference compresses ->
decompress; identity.
"""
import ctypes as C
import random
import subprocess

import pytest

from conftest import golden_bytes, orc_decode, ref_conduit_decode, gen_bytes, sz, dict_compress, dict_decompress

pytestmark = pytest.mark.gpu


def payloads(manifest, gen):
    out = {}
    for k, e in manifest.items():
        if not k.startswith("_"):
            out[k] = golden_bytes(e["raw"])[:e["size"]]
    # multi-batch (B2LZ4_BATCH_MB=8 in conftest): log-like + zeros + random, odd size
    out["big"] = gen_bytes(gen, "loglike", 19 * 1024 * 1024 + 4321) + bytes(700001) + gen_bytes(gen, "random", 300007)
    return out


def b200_conduit_compress(replay, data, chunk=32768, yi=False):
    cap = len(data) + len(data) // 50 + (96 << 20)
    o = C.create_string_buffer(cap)
    if yi:
        r = replay.ref_conduit_compress_yi(data, len(data), o, cap, chunk)
    else:
        r = replay.ref_conduit_compress(data, len(data), o, cap, chunk, 0)
    assert r < (1 << 62), hex(r)
    return o.raw[:r]


def test_compress_conduit_replay_decodes_everywhere(replay, oracle, ref, lz4_cli, manifest, gen, tmp_path):
    """test/Main.hs:60-85: what `compress` emits must decode with the stock CLI (and the reference library)."""
    for name, d in payloads(manifest, gen).items():
        fr = b200_conduit_compress(replay, d)
        assert fr[:4] == b"\x04\x22\x4d\x18"
        got, consumed = orc_decode(oracle, fr, len(d) + 64)
        assert got == d and consumed == len(fr), name
        if ref is not None:
            got, consumed = ref_conduit_decode(ref, fr, len(d) + 64, 16384)
            assert got == d and consumed == len(fr), name
        if lz4_cli is not None and name in ("hellos", "numbers", "big", "empty"):
            p = tmp_path / (name + ".lz4")
            p.write_bytes(fr)
            assert subprocess.run([lz4_cli, "-dc", str(p)], capture_output=True, check=True).stdout == d, name


def test_compress_is_chunking_independent_and_yield_immediately_agrees(replay, gen):
    """the frame does not depend on how upstream cut the ByteStrings; compressYieldImmediately
    (Conduit.hsc:364-425) produces the same bytes as compress."""
    d = gen_bytes(gen, "loglike", 9 * 1024 * 1024 + 77)
    a = b200_conduit_compress(replay, d, 32768)
    assert a == b200_conduit_compress(replay, d, 1000)
    assert a == b200_conduit_compress(replay, d, 0)
    assert a == b200_conduit_compress(replay, d, 3 << 20, yi=True)


def test_decompress_conduit_replay_on_reference_frames(replay, manifest, gen, ref):
    """test/Main.hs:87-112 + SURVEY §8: frames made by the reference conduit (linked 64 KB blocks), the stock CLI
    (independent, checksums) and the library with flushes decode bit-exactly through LZ4F_getFrameInfo +
    LZ4F_decompress with a 16 KB destination, for several upstream chunk sizes; the frame is consumed exactly."""
    for name in [k for k in manifest if not k.startswith("_")]:
        d = golden_bytes(manifest[name]["raw"])[:manifest[name]["size"]]
        for kind, fr in manifest[name]["frames"].items():
            b = golden_bytes(fr["file"])
            for chunk in (0, 32768, 1000):
                got, consumed = ref_conduit_decode(replay, b, len(d) + 64, chunk)
                assert got == d, (name, kind, chunk, got if isinstance(got, int) else len(got))
                assert consumed == len(b), (name, kind, chunk)
    if ref is not None:                                     # multi-batch linked frame: history crosses GPU batches
        d = payloads(manifest, gen)["big"]
        cap = len(d) + len(d) // 100 + 100000
        o = C.create_string_buffer(cap)
        r = ref.ref_conduit_compress(d, len(d), o, cap, 32768, 0)
        fr = o.raw[:r]
        for chunk in (0, 65536, 12345):
            got, consumed = ref_conduit_decode(replay, fr, len(d) + 64, chunk)
            assert got == d and consumed == len(fr), chunk
        for args in [(7, 1, 1, 1, 1, 1, 4 << 20), (4, 0, 1, 0, 0, 0, 50000), (5, 1, 0, 1, 0, 1, 100000)]:
            r = ref.ref_frame_compress(d, len(d), o, cap, *args)
            fr = o.raw[:r]
            got, consumed = ref_conduit_decode(replay, fr, len(d) + 64, 100000)
            assert got == d and consumed == len(fr), args


def test_identity_and_trailing_bytes(replay, gen, ref):
    """test/Main.hs:114-119 (QuickCheck identity) + Conduit.hsc:693-695 (stop at the first EndMark)."""
    rnd = random.Random(11)
    for _ in range(12):
        n = rnd.choice([0, 1, 12, 13, 100, 5000, 65535, 65536, 65537, 300000])
        d = bytes(rnd.getrandbits(8) if rnd.random() < 0.3 else 65 + (i % 7) for i in range(n))
        fr = b200_conduit_compress(replay, d, 4096)
        # the frame is one upstream ByteString, the trailing bytes the next one: the conduit stops after the first
        got, consumed = ref_conduit_decode(replay, fr + b"tail", len(d) + 64, len(fr) - 7)
        assert got == d and consumed == len(fr), n


def test_trailing_bytes_inside_the_last_chunk_behave_like_the_reference(replay, ref, manifest):
    """When the frame ends INSIDE an upstream ByteString the Haskell code feeds the rest to LZ4F_decompress again
    (Conduit.hsc:682-683): whatever the reference library answers there, we must answer too."""
    if ref is None:
        pytest.skip("oracle/_ref not present")
    d = golden_bytes(manifest["hellos"]["raw"])[:manifest["hellos"]["size"]]
    b = golden_bytes(manifest["hellos"]["frames"]["conduit"]["file"])
    for tail in (b"tail", b, b"\x00" * 9):
        want = ref_conduit_decode(ref, b + tail, 2 * len(d) + 64, 0)
        got = ref_conduit_decode(replay, b + tail, 2 * len(d) + 64, 0)
        assert got == want, (tail[:8], got if isinstance(got[0], int) else len(got[0]), want if isinstance(want[0], int) else len(want[0]))


def test_preferences_checksums_and_content_size(replay, oracle, ref, gen):
    """CTypes.hsc:155-232 Preferences reach the encoder: block checksum, content checksum, content size;
    a wrong content size is ERROR_frameSize_wrong (lz4frame.c:1012-1015)."""
    d = gen_bytes(gen, "loglike", 5 * 1024 * 1024 + 3) + gen_bytes(gen, "random", 70000)
    cap = len(d) + (96 << 20)
    o = C.create_string_buffer(cap)
    for args in [(7, 1, 1, 1, 1, 0, 1 << 20), (4, 1, 1, 0, 0, 1, 65536), (6, 0, 0, 1, 1, 1, 333333), (0, 0, 0, 0, 0, 0, 16384)]:
        r = replay.ref_frame_compress(d, len(d), o, cap, *args)
        assert r < (1 << 62), (args, hex(r))
        fr = o.raw[:r]
        got, consumed = orc_decode(oracle, fr, len(d) + 64)
        assert got == d and consumed == len(fr), args
        if ref is not None:
            got, consumed = ref_conduit_decode(ref, fr, len(d) + 64, 16384)
            assert got == d and consumed == len(fr), args
        got, consumed = ref_conduit_decode(replay, fr, len(d) + 64, 50000)
        assert got == d and consumed == len(fr), args
        if args[2]:                                          # flip one content byte region -> content checksum must trip
            bad = bytearray(fr); bad[-2] ^= 0x40
            got, _ = ref_conduit_decode(replay, bytes(bad), len(d) + 64, 0)
            assert got == -18, got                           # ERROR_contentChecksum_invalid


def test_truncated_stream_reports_no_endmark(replay, manifest):
    """Conduit.hsc:689: input that ends before the EndMark is an error, not a short result."""
    b = golden_bytes(manifest["mixed166k"]["frames"]["conduit"]["file"])
    got, _ = ref_conduit_decode(replay, b[:-4], 400000, 4096)
    assert got == -102, got                                  # REPLAY_ERR_NO_ENDMARK (the conduit's throwString)
    got, _ = ref_conduit_decode(replay, b[:3], 400000, 0)
    assert got == -101, got                                  # "not enough bytes for header" (Conduit.hsc:615-616)


# ------------------------------------------------------------------ linked-block encoding and dictionaries
def test_linked_block_encoding_ratio_and_interop(pkg, oracle, ref, lz4_cli, datagen, gen):
    """blockMode = linked honoured (FLG 0x40): every decoder takes the frames; at 64 KB blocks — the reference conduit's own
    shape, BASELINE config 1 — the size is within the 2 % of the reference's default that independent 64 KB blocks miss."""
    assert ref is not None and datagen
    text = subprocess.run([datagen, "-g67108864", "-s1"], capture_output=True, check=True).stdout
    shapes = {"datagen 64 MiB": text, "log-like 32 MiB": gen_bytes(gen, "loglike", 32 << 20),
              "mix": bytes(300000) + gen_bytes(gen, "random", 200000, 2) + gen_bytes(gen, "textlike", 900000, 4), "tiny": b"ab" * 7}
    with pkg.Codec(0, use_index=0) as c:
        for bid in (4, 7):
            c.set_prefs(block_size_id=bid, sub_block_log2=16 if bid == 4 else 17, block_linked=1, write_index=0, content_checksum=1)
            for what, d in shapes.items():
                fr = c.compress_bytes(d)
                assert fr[4] & 0x20 == 0
                got, used = orc_decode(oracle, fr, len(d) + 16)
                assert got == d and used == len(fr), (bid, what)
                assert c.decompress_bytes(fr)[0] == d, (bid, what)
                got, used = ref_conduit_decode(ref, fr, len(d) + 64)
                assert got == d, (bid, what)
                if lz4_cli and len(d) < (8 << 20):
                    assert subprocess.run([lz4_cli, "-d", "-c"], input=fr, capture_output=True).stdout == d
                if len(d) > (1 << 20) and what != "mix":
                    cap = len(d) + len(d) // 100 + 100000
                    o = C.create_string_buffer(cap)
                    r = ref.ref_conduit_compress(d, len(d), o, cap, 0, 0)
                    print(f"linked {65536 << (2 * (bid - 4))} B blocks, {what}: ours {len(fr)} B, reference default {r} B ({100.0 * (len(fr) / r - 1):+.2f} %)")
                    assert len(fr) <= r * 1.02, (bid, what, len(fr), r)


def test_dictionary_frames_both_ways(replay, ref, lz4_cli, manifest, gen, tmp_path):
    """LZ4F_createCDict / LZ4F_compressBegin_usingCDict / LZ4F_decompress_usingDict on the product's symbols: reference-made
    dictionary frames (goldens) decode here; frames made here decode with the reference library and `lz4 -d -D`."""
    dman = manifest["_dict"]
    dic = golden_bytes(dman["file"])
    for name, fr in dman["frames"].items():
        raw = golden_bytes(manifest[name]["raw"])[:manifest[name]["size"]]
        f = golden_bytes(fr["file"])
        assert dict_decompress(replay, f, dic, len(raw) + 64, 7000) == raw, name
        if name in ("datagen200k", "mixed166k"):
            assert dict_decompress(replay, f, b"", len(raw) + 64) != raw, name
    dfile = tmp_path / "dict"
    dfile.write_bytes(dic)
    big = dic[100:30000] + gen_bytes(gen, "loglike", 20 << 20)              # several batches (B2LZ4_BATCH_MB=8)
    for d in (b"", b"x", dic[5000:5300], dic[100:9000] + gen_bytes(gen, "loglike", 100000), gen_bytes(gen, "textlike", 300000, 3), big):
        for did in (0, 0xABCD1234):
            ours = dict_compress(replay, d, dic, 4, 0, 1, did)
            assert ours[4] & 0x20 == 0 and bool(ours[4] & 1) == bool(did)
            assert dict_decompress(replay, ours, dic, len(d) + 64, 99999) == d
            if ref is not None:
                assert dict_decompress(ref, ours, dic, len(d) + 64, 100000) == d
                indep = dict_compress(ref, d, dic, 4, 1, 1, did)                  # independent blocks + dictionary (what `lz4 -D` writes for one block)
                assert dict_decompress(replay, indep, dic, len(d) + 64, 4096) == d
                theirs = dict_compress(ref, d, dic, 4, 0, 1, did)
                assert ours[:11 if did else 7] == theirs[:11 if did else 7]
                assert dict_decompress(replay, theirs, dic, len(d) + 64) == d
                if 200 < len(d) < 10000:
                    assert len(ours) < len(dict_compress(replay, d, b"", 4, 0, 1, did))
            if lz4_cli and len(d) < (1 << 20):
                r = subprocess.run([lz4_cli, "-d", "-c", "-D", str(dfile)], input=ours, capture_output=True)
                assert r.returncode == 0 and r.stdout == d
