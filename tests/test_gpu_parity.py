"""GPU parity tests (run on the B200 box): every call goes through the C ABI (libb200lz4.so).

  decoder : bit-exact against the oracle / the reference on frames the reference and the stock CLI made
  encoder : every frame we emit decodes bit-exactly with the oracle, the reference conduit replay and
            the stock `lz4 -d`; size within tolerance of the reference's default preferences
  XXH32   : equal to the oracle on every length class
"""
import ctypes as C
import os
import random
import subprocess

import pytest

from conftest import golden_bytes, orc_decode, ref_conduit_decode, gen_bytes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def codec(pkg):
    c = pkg.Codec(0)
    yield c
    c.close()


def payload(man, name):
    e = man[name]
    return golden_bytes(e["raw"])[:e["size"]]


def names(man):
    return [k for k in man if not k.startswith("_")]


# ------------------------------------------------------------------ XXH32
def test_xxh32_matches_oracle(codec, oracle, gen, manifest):
    kat = manifest["_xxh32_kat"]
    assert codec.xxh32(b"") == kat[""] == 0x02CC5D05
    assert codec.xxh32(b"abc") == kat["abc"] == 0x32D153FF
    d = gen_bytes(gen, "random", 200003)
    for n in (1, 3, 4, 15, 16, 17, 31, 32, 33, 4095, 16384, 16385, 32768 + 5, 100000, 200003):
        assert codec.xxh32(d[:n]) == oracle.orc_xxh32(d, n, 0), n
    for off in (1, 2, 3, 5):                                      # misaligned device source
        assert codec.xxh32(d[off:off + 70001]) == oracle.orc_xxh32(d[off:], 70001, 0)


# ------------------------------------------------------------------ decoder on reference-made frames
def test_decode_every_golden_frame(codec, manifest):
    """Frames made by the reference conduit replay (linked 64 KB), the stock CLI (independent, checksums)
    and the reference library with flushes: output must be bit-exact and the whole frame consumed."""
    for name in names(manifest):
        d = payload(manifest, name)
        for kind, fr in manifest[name]["frames"].items():
            b = golden_bytes(fr["file"])
            out, consumed = codec.decompress_bytes(b, max_out=len(d) + 64)
            assert out == d, (name, kind, len(out), len(d))
            assert consumed == len(b), (name, kind)


def test_decode_exact_capacity_and_too_small(codec, manifest, pkg):
    d = payload(manifest, "numbers")
    b = golden_bytes(manifest["numbers"]["frames"]["cli_B4_BX"]["file"])
    out, _ = codec.decompress_bytes(b, max_out=len(d))
    assert out == d
    with pytest.raises(pkg.Lz4FrameError) as ei:
        codec.decompress_bytes(b, max_out=len(d) - 1)
    assert ei.value.name == "ERROR_dstMaxSize_tooSmall"


def test_decode_matches_oracle_on_reference_frames_live(codec, oracle, ref, gen):
    """Bigger frames made on the spot by the unmodified reference library (oracle/_ref travels)."""
    if ref is None:
        pytest.skip("oracle/_ref not present")
    d = gen_bytes(gen, "loglike", 9 * 1024 * 1024 + 12345) + bytes(300000) + gen_bytes(gen, "random", 300000)
    cap = len(d) + len(d) // 100 + 100000
    o = C.create_string_buffer(cap)
    r = ref.ref_conduit_compress(d, len(d), o, cap, 32768, 0)          # reference conduit: linked 64 KB blocks
    assert not ref.LZ4F_isError(r)
    frames = {"conduit": o.raw[:r]}
    for (bsid, indep, cc, bc, cs, chunk) in [(7, 1, 1, 0, 0, 4 << 20), (4, 1, 1, 1, 0, 65536), (6, 0, 1, 0, 1, 1 << 20),
                                             (5, 1, 0, 1, 0, 100000), (4, 0, 0, 0, 0, 50000)]:
        r = ref.ref_frame_compress(d, len(d), o, cap, bsid, indep, cc, bc, cs, 1, chunk)
        assert not ref.LZ4F_isError(r)
        frames[(bsid, indep, cc, bc, cs, chunk)] = o.raw[:r]
    for k, fr in frames.items():
        want, consumed_want = orc_decode(oracle, fr, len(d) + 16)
        assert want == d
        got, consumed = codec.decompress_bytes(fr, max_out=len(d) + 16)
        assert got == want, k
        assert consumed == consumed_want == len(fr), k


def test_decode_stops_at_first_frame(codec, manifest):
    """The reference conduit stops at the first EndMark and leaves the rest upstream (Conduit.hsc:693-695)."""
    d = payload(manifest, "hellos")
    b = golden_bytes(manifest["hellos"]["frames"]["cli"]["file"])
    out, consumed = codec.decompress_bytes(b + b"trailing garbage" + b, max_out=len(d) + 64)
    assert out == d and consumed == len(b)


def test_decode_rejects_corruption(codec, oracle, manifest, pkg):
    """lz4/tests/frametest.c:318-331, 1044-1073: corrupted frames must error (never crash); accept/reject and
    bytes must agree with the oracle."""
    rnd = random.Random(5)
    d = payload(manifest, "mixed166k")
    for kind in ("cli_B4_BX", "conduit", "cli"):
        good = golden_bytes(manifest["mixed166k"]["frames"][kind]["file"])
        for trial in range(40):
            b = bytearray(good)
            if trial % 4 == 3:
                b = b[:rnd.randrange(7, len(b))]
            else:
                for _ in range(rnd.randint(1, 3)):
                    b[rnd.randrange(7, len(b))] ^= 1 << rnd.randrange(8)
            b = bytes(b)
            want, _ = orc_decode(oracle, b, len(d) + 64)
            try:
                got, _ = codec.decompress_bytes(b, max_out=len(d) + 64)
            except pkg.Lz4FrameError:
                got = None
            if isinstance(want, int):
                assert got is None, (kind, trial, want)
            else:
                assert got == want, (kind, trial)


def test_decode_fuzz_never_writes_outside_the_destination(codec, oracle, manifest, pkg, gen):
    """Bounds check of our own (compute-sanitizer is not available on the GPU pool): corrupted and truncated frames are
    decoded DEVICE-resident into a destination of exactly the payload's size that sits between two guard zones, and
    the compressed input has a guard behind it as well.  Whatever the verdict, the guards must be untouched, and
    verdict and bytes must agree with the oracle (lz4/tests/frametest.c:1044-1073, fuzzer.c:1040-1053)."""
    import torch
    rnd = random.Random(2026)
    G = 1 << 16
    cases = [(payload(manifest, "mixed166k"), golden_bytes(manifest["mixed166k"]["frames"][k]["file"])) for k in ("cli_B4_BX", "conduit", "cli")]
    d2 = gen_bytes(gen, "loglike", 700000)
    for prefs in (dict(block_size_id=4, write_index=1), dict(block_size_id=5, sub_block_log2=16, write_index=0, block_checksum=1)):
        codec.set_prefs(**{**dict(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18,
                                  write_index=1, use_index=1), **prefs})
        cases.append((d2, codec.compress_bytes(d2)))
    codec.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18, write_index=1, use_index=1)
    for d, good in cases:
        n = len(d)
        for trial in range(30):
            b = bytearray(good)
            kind = trial % 5
            if kind == 4:
                b = b[:rnd.randrange(7, len(b))]
            elif kind == 3:                                  # length fields: tokens / extension bytes forced to 255
                for _ in range(rnd.randint(1, 4)):
                    b[rnd.randrange(7, len(b))] = 0xFF
            else:
                for _ in range(rnd.randint(1, 3)):
                    b[rnd.randrange(7, len(b))] ^= 1 << rnd.randrange(8)
            b = bytes(b)
            want, _ = orc_decode(oracle, b, n)
            src = torch.full((len(b) + G,), 0xA5, dtype=torch.uint8, device="cuda")
            src[:len(b)] = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
            dst = torch.full((n + 2 * G,), 0xEE, dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            try:
                r, _ = codec.decompress_device(src.data_ptr(), len(b), dst.data_ptr() + G, n)
                got = bytes(dst[G:G + r].cpu().numpy())
            except pkg.Lz4FrameError:
                got = None
            torch.cuda.synchronize()
            assert int(dst[:G].min()) == 0xEE and int(dst[G + n:].min()) == 0xEE, ("guard overwritten", trial)
            assert int(dst[:G].max()) == 0xEE and int(dst[G + n:].max()) == 0xEE, ("guard overwritten", trial)
            assert int(src[len(b):].min()) == 0xA5 and int(src[len(b):].max()) == 0xA5
            if isinstance(want, int):
                assert got is None, (trial, want)
            else:
                assert got == want, trial


def test_corrupt_index_never_fails_a_valid_frame(codec, gen):
    """The B2IX index travels in a skippable frame BEHIND the frame (lz4_Frame_format.md:298-335): a stock decoder ignores
    it, so whatever is wrong with it must not change the verdict on the frame.  An entry that passes the structural
    checks but points one byte off makes the indexed decode fail; the engine then decodes without the index — in the
    single-call device path and in the chunked host-buffer path (frame > one pipeline chunk)."""
    d = gen_bytes(gen, "loglike", 96 << 20)
    codec.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=17, write_index=1, use_index=1)
    fr = bytearray(codec.compress_bytes(d))
    at = len(fr) - 8 - 8                                         # token offset of the last chain of the last block
    tok = int.from_bytes(fr[at:at + 4], "little")
    assert tok > 0
    fr[at:at + 4] = (tok + 1).to_bytes(4, "little")
    out, consumed = codec.decompress_bytes(bytes(fr), max_out=len(d))          # host buffers: chunked pipeline, index slices
    assert out == d
    small = gen_bytes(gen, "loglike", 9 << 20)
    fr = bytearray(codec.compress_bytes(small))
    at = len(fr) - 8 - 8
    fr[at:at + 4] = (int.from_bytes(fr[at:at + 4], "little") + 1).to_bytes(4, "little")
    out, _ = codec.decompress_bytes(bytes(fr), max_out=len(small))             # one call on the device
    assert out == small and codec.stats().retries >= 1


def test_decode_malformed_blocks(codec, pkg):
    # lz4/tests/fuzzer.c:1040-1053: the 17-byte block that must be rejected, wrapped in a frame
    hdr = bytes.fromhex("04224d18604082")
    for blk in (bytes([0xEE]) + bytes(14) + bytes([0x0E, 0x00]), b"\xff", b"\x10", b"\x00\x00"):
        fr = hdr + len(blk).to_bytes(4, "little") + blk + bytes(4)
        with pytest.raises(pkg.Lz4FrameError):
            codec.decompress_bytes(fr, max_out=4096)
    # a frame whose block size word exceeds the maximum (lz4frame.c:1489-1490)
    with pytest.raises(pkg.Lz4FrameError) as ei:
        codec.decompress_bytes(hdr + (70000).to_bytes(4, "little") + bytes(70004), max_out=100000)
    assert ei.value.name == "ERROR_maxBlockSize_invalid"
    # truncated: no EndMark (Conduit.hsc:689)
    with pytest.raises(pkg.Lz4FrameError):
        codec.decompress_bytes(hdr + (5).to_bytes(4, "little") + b"\x40abcd", max_out=4096)
    # empty frame is fine
    out, consumed = codec.decompress_bytes(hdr + bytes(4), max_out=16)
    assert out == b"" and consumed == 11


# ------------------------------------------------------------------ encoder
PREF_SETS = [dict(), dict(write_index=0), dict(block_size_id=4), dict(block_size_id=5, block_checksum=1), dict(block_size_id=6, content_checksum=1),
             dict(block_size_id=7, block_checksum=1, content_checksum=1, content_size=1),
             dict(block_size_id=7, sub_block_log2=16), dict(block_size_id=4, write_index=0)]


def roundtrip_everywhere(codec, oracle, ref, lz4_cli, d, prefs, tmp_path=None):
    codec.set_prefs(**{**dict(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0,
                              sub_block_log2=18, write_index=1, use_index=1), **prefs})
    fr = codec.compress_bytes(d)
    # (1) our decoder, with and without the index
    out, consumed = codec.decompress_bytes(fr, max_out=len(d) + 16)
    assert out == d and consumed == len(fr)
    codec.set_prefs(use_index=0)
    out, consumed2 = codec.decompress_bytes(fr, max_out=len(d) + 16)
    codec.set_prefs(use_index=1)
    assert out == d and consumed2 <= consumed
    # (2) the oracle
    want, oc = orc_decode(oracle, fr, len(d) + 16)
    assert want == d and oc == consumed2
    # (3) the reference conduit replay (16 KB output slices, odd input chunking)
    if ref is not None:
        got, rc = ref_conduit_decode(ref, fr, len(d) + 16, chunk=10007)
        # within one ByteString the reference keeps feeding LZ4F_decompress after the EndMark (Conduit.hsc:681-684),
        # so a trailing skippable index frame in the same chunk is skipped and counted as consumed
        assert got == d and rc in (consumed2, consumed)
    # (4) the stock CLI (skips our trailing skippable index frame, lz4io.c:1094-1135)
    if lz4_cli is not None:
        if codec.prefs.write_index:          # a file: the v1.9.2 CLI skips skippable frames with fseek (lz4io.c:1124-1133)
            import tempfile
            with tempfile.NamedTemporaryFile(suffix=".lz4") as f:
                f.write(fr); f.flush()
                r = subprocess.run([lz4_cli, "-d", "-c", f.name], capture_output=True)
        else:                                # a pipe, like test/Main.hs:27-31
            r = subprocess.run([lz4_cli, "-d", "-c"], input=fr, capture_output=True)
        assert r.returncode == 0, r.stderr
        assert r.stdout == d
    return fr


def test_encode_reference_test_payloads(codec, oracle, ref, lz4_cli, manifest):
    """The payload shapes of test/Main.hs:60-112, every preference set."""
    for name in names(manifest):
        d = payload(manifest, name)
        for prefs in PREF_SETS:
            roundtrip_everywhere(codec, oracle, ref, lz4_cli, d, prefs)


def test_encode_block_boundaries(codec, oracle, ref, lz4_cli, gen):
    base = gen_bytes(gen, "loglike", (4 << 20) + 70000)
    for n in (1, 12, 13, 65535, 65536, 65537, 262143, 262144, 262145, 1200007, (4 << 20) - 1, 4 << 20, (4 << 20) + 1):
        for prefs in (dict(), dict(block_size_id=4, block_checksum=1, content_checksum=1)):
            roundtrip_everywhere(codec, oracle, ref, None, base[:n], prefs)
    roundtrip_everywhere(codec, oracle, ref, lz4_cli, base, dict())


def test_encode_edge_mix(codec, oracle, ref, lz4_cli, gen):
    """Config 5 shapes at test size: incompressible (stored blocks) and all-zero (long overlapping matches)."""
    rnd = gen_bytes(gen, "random", 9 << 20, seed=7)
    fr = roundtrip_everywhere(codec, oracle, ref, lz4_cli, rnd, dict())
    assert len(fr) <= len(rnd) + 3 * 8 + 7 + 4 + 200            # stored: 4 bytes per block + header + EndMark + index
    assert fr[7:11] == bytes.fromhex("00004080")                 # header word of a stored 4 MB block
    z = bytes(9 << 20)
    fr = roundtrip_everywhere(codec, oracle, ref, lz4_cli, z, dict())
    assert len(fr) < (9 << 20) // 200
    roundtrip_everywhere(codec, oracle, ref, lz4_cli, z[:100000] + rnd[:100000] + z[:70000], dict(block_size_id=4))
    # short periods 1..40: overlapping match copies with every small offset
    pat = b"".join(bytes(range(65, 65 + p)) * (3000 // p) for p in range(1, 41))
    roundtrip_everywhere(codec, oracle, ref, lz4_cli, pat, dict(block_size_id=4))


def test_encode_ratio_vs_reference_default(codec, ref, gen, datagen):
    """north_star: compressed size <= 2 % larger than the reference's default LZ4F preferences on the same
    input.  Tolerance stated here: 2 % at 4 MB blocks, for the log-like config-2 shape and for the reference's own
    text generator (`datagen -s1`)."""
    if ref is None:
        pytest.skip("oracle/_ref not present")
    d = gen_bytes(gen, "loglike", 32 << 20)
    cap = len(d) + len(d) // 100 + 100000
    o = C.create_string_buffer(cap)
    r_ref = ref.ref_conduit_compress(d, len(d), o, cap, 0, 0)
    codec.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18, write_index=0)
    ours = len(codec.compress_bytes(d))
    codec.set_prefs(write_index=1)
    print(f"log-like 32 MiB: reference default {r_ref} B, ours {ours} B ({100.0 * (ours / r_ref - 1):+.2f} %)")
    assert ours <= r_ref * 1.02
    if datagen:
        t = subprocess.run([datagen, "-g16777216", "-s1"], capture_output=True, check=True).stdout
        r_ref = ref.ref_conduit_compress(t, len(t), o, cap, 0, 0)
        ours = len(codec.compress_bytes(t))
        print(f"datagen 16 MiB: reference default {r_ref} B, ours {ours} B ({100.0 * (ours / r_ref - 1):+.2f} %)")
        assert ours <= r_ref * 1.02          # north_star tolerance at the stated 4 MB block size


def test_block_checksums_of_every_block_size_both_ways(codec, lz4_cli, gen):
    """blockChecksumFlag (lz4frame.c:756-759, 1604-1611; `lz4 -BX`) at every block size: small blocks are hashed by a warp each,
    blocks of 256 KB and more by a CTA each (xxh32_cta.cuh).  Our frames must pass `lz4 -d` (which verifies every block), the
    stock CLI's frames must decode here, and a flipped bit in one must be caught — by the block checksum, not later."""
    if lz4_cli is None:
        pytest.skip("oracle/_ref/lz4 not present")
    d = gen_bytes(gen, "loglike", (9 << 20) + 4321) + os.urandom(70001)
    for bid in (4, 5, 6, 7):
        codec.set_prefs(block_size_id=bid, sub_block_log2=16, block_checksum=1, content_checksum=0, content_size=0, write_index=0, use_index=1)
        fr = codec.compress_bytes(d)
        r = subprocess.run([lz4_cli, "-d", "-c"], input=fr, capture_output=True)
        assert r.returncode == 0 and r.stdout == d, (bid, r.stderr[:200])
        f2 = subprocess.run([lz4_cli, "-c", "-BX", f"-B{bid}", "--no-frame-crc"], input=d, capture_output=True, check=True).stdout
        out, used = codec.decompress_bytes(f2, max_out=len(d) + 16)
        assert out == d and used == len(f2), bid
        bad = bytearray(f2)
        bad[len(bad) // 3] ^= 0x10
        with pytest.raises(Exception) as ei:
            codec.decompress_bytes(bytes(bad), max_out=len(d) + 16)
        assert "blockChecksum_invalid" in str(ei.value) or "decompressionFailed" in str(ei.value) or "maxBlockSize" in str(ei.value), (bid, str(ei.value))
    codec.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18, write_index=1, use_index=1)


def test_default_prefs_emit_plain_frames(pkg, lz4_cli, gen):
    """Drop-in default: no trailing index frame, `lz4 -d` on a pipe accepts the output (test/Main.hs:60-85)."""
    with pkg.Codec(0) as c:
        assert c.prefs.write_index == 0 and c.prefs.block_size_id == 5
        d = gen_bytes(gen, "loglike", 3 << 20)
        fr = c.compress_bytes(d)
        assert fr[:7].hex() == "04224d186050fb" and fr[-4:] == bytes(4)
        if lz4_cli:
            r = subprocess.run([lz4_cli, "-d"], input=fr, capture_output=True)
            assert r.returncode == 0 and r.stdout == d


def test_compress_is_deterministic(codec, gen):
    d = gen_bytes(gen, "loglike", 5 << 20)
    assert codec.compress_bytes(d) == codec.compress_bytes(d)


def test_conduit_generators(pkg, codec, manifest):
    """compress .| decompress == identity (test/Main.hs:114-119, QuickCheck identity) through the conduit mirror."""
    rnd = random.Random(11)
    for _ in range(20):
        n = rnd.randrange(0, 10000)
        s = bytes(rnd.choice(b"abcdefgh \n") for _ in range(n))
        chunks = pkg.bs_chunks_of(rnd.randrange(1, 5000), s)
        fr = b"".join(pkg.compress(chunks, codec))
        assert b"".join(pkg.decompress(pkg.bs_chunks_of(777, fr), codec)) == s
    with pytest.raises(pkg.Lz4FrameError):
        list(pkg.decompress([golden_bytes(manifest["numbers"]["frames"]["conduit"]["file"])[:-6]], codec))


# ------------------------------------------------------------------ device-resident, full size properties
def test_device_resident_roundtrip_256MiB(codec, gen):
    """Size-independent property at a large size: decode(encode(x)) == x, checked on the device."""
    import torch
    n = 256 << 20
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    gen.b2gen_loglike(host.data_ptr(), n, 0x10C5)
    x = host.cuda()
    cap = 300 << 20
    comp = torch.empty(cap, dtype=torch.uint8, device="cuda")
    back = torch.zeros(n + 4096, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    codec.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18, write_index=1, use_index=1)
    c = codec.compress_device(x.data_ptr(), n, comp.data_ptr(), cap)
    st = codec.stats()
    assert st.n_blocks == 64 and st.indexed == 1
    r, consumed = codec.decompress_device(comp.data_ptr(), c, back.data_ptr(), n + 4096)
    assert r == n and consumed == c
    assert torch.equal(back[:n], x)
    assert int(back[n:].max()) == 0                               # canary: nothing written past the output
    assert 2.0 < n / c < 4.0
