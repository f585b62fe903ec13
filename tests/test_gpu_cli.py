"""b200lz4c — the command-line twin of the reference's haskell-lz4c (app/Main.hs) — against the stock lz4 CLI.

These are the reference's own hspec cases (test/Main.hs:60-112: compress with the conduit -> `lz4 -d`, `lz4` -> decompress
with the conduit, for "hellohellohellohello", 100000 integers, 100000 strings, 1 MB of one byte, and the > 32-bit case)
run through the product binary, plus the SURVEY §8(f) rows the CLI adds: Preferences flags and `lz4 -d` style handling
of concatenated / skippable / legacy frames.  The stock CLI (oracle/_ref/lz4) is only ever the checker."""
import os
import struct
import subprocess

import pytest

from conftest import ROOT, gen_bytes

pytestmark = pytest.mark.gpu
CLI = os.path.join(ROOT, "lz4-frame-conduit_b200", "b200lz4c")


def run(cmd, data=b"", ok=True):
    p = subprocess.run(cmd, input=data, capture_output=True, timeout=600)
    if ok:
        assert p.returncode == 0, p.stderr.decode()[-500:]
    return p


def payloads():
    return {
        "simple": b"hellohellohellohello",                                        # test/Main.hs:62
        "integers": b"".join(str(i).encode() for i in range(1, 100001)),          # :67
        "strings": b"hello" * 100000,                                             # :72
        "1MB": bytes([42]) * 100000,                                              # :77
        "empty": b"",
    }


@pytest.fixture(scope="module", autouse=True)
def _need_cli():
    if not os.path.exists(CLI):
        subprocess.run(["make", "-C", os.path.join(ROOT, "lz4-frame-conduit_b200", "csrc")], check=True, capture_output=True)


def test_cli_compress_then_stock_lz4_decodes(lz4_cli):
    """runCompressToLZ4 (test/Main.hs:36-40): our compress | lz4 -d == input."""
    assert lz4_cli
    for name, data in payloads().items():
        frame = run([CLI], data).stdout
        assert frame[:4] == b"\x04\x22\x4d\x18"
        assert run([lz4_cli, "-d", "-c"], frame).stdout == data, name


def test_cli_decompresses_stock_lz4_output(lz4_cli):
    """runLZ4ToDecompress (test/Main.hs:42-46): lz4 | our decompress == input — default flags and the config-4 flag sets."""
    assert lz4_cli
    for flags in ([], ["-B4", "-BX"], ["-B5", "-BD"], ["-B6", "--no-frame-crc"], ["-B7", "-BX", "--content-size"]):
        for name, data in payloads().items():
            frame = run([lz4_cli, "-c"] + flags, data).stdout
            assert run([CLI, "-d"], frame).stdout == data, (name, flags)


def test_cli_identity_with_files_and_preferences(tmp_path, lz4_cli, gen):
    data = gen_bytes(gen, "loglike", (9 << 20) + 777)
    src, mid, back = tmp_path / "in.raw", tmp_path / "in.lz4", tmp_path / "out.raw"
    src.write_bytes(data)
    for flags, hdr in (([], b"\x60\x50"), (["-B4", "--block-checksum", "--content-checksum"], b"\x74\x40"),
                       (["-B6", "--content-checksum"], b"\x64\x60")):
        run([CLI, str(src), str(mid)] + flags)
        frame = mid.read_bytes()
        assert frame[4:6] == hdr, flags                      # FLG / BD bytes (lz4_Frame_format.md:136-145)
        run([CLI, "-d", str(mid), str(back)])
        assert back.read_bytes() == data
        assert run([lz4_cli, "-d", "-c", str(mid)]).stdout == data


def test_cli_all_frames_like_lz4_d(lz4_cli, gen, tmp_path):
    """Concatenated frames, a skippable frame in between, and a legacy frame: --all-frames decodes what `lz4 -d` decodes
    (lz4io.c:1094-1135); without it only the first frame is decoded, like the reference conduit (Conduit.hsc:598-701)."""
    a = gen_bytes(gen, "loglike", 3 << 20, 1)
    b = gen_bytes(gen, "textlike", (17 << 20) + 5, 2)        # legacy: 8 MB blocks -> 3 blocks
    c = gen_bytes(gen, "random", 70000, 3) + bytes(50000)
    fa = run([lz4_cli, "-c"], a).stdout
    fb = run([lz4_cli, "-c", "-l"], b).stdout
    assert fb[:4] == struct.pack("<I", 0x184C2102)
    fc = run([CLI, "-B5"], c).stdout
    skip = struct.pack("<II", 0x184D2A53, 11) + b"hello world"
    stream = fa + skip + fb + skip + fc
    f = tmp_path / "multi.lz4"                              # the stock CLI skips skippable frames with fseek: files only (lz4io.c:1124-1133)
    f.write_bytes(stream)
    want = run([lz4_cli, "-d", "-c", str(f)]).stdout
    assert want == a + b + c
    assert run([CLI, "-d", "--all-frames"], stream).stdout == want
    assert run([CLI, "-d"], stream).stdout == a              # first frame only
    # a legacy frame alone, and garbage after a frame
    assert run([CLI, "-d", "--all-frames"], fb).stdout == b
    p = run([CLI, "-d", "--all-frames"], fa + b"garbage!")
    assert p.stdout == a and b"undecodable" in p.stderr


def test_cli_legacy_low_ratio_blocks_over_4mib(lz4_cli):
    """`lz4 -l` blocks are 8 MB and, at a ratio near 1.3, their COMPRESSED size exceeds 4 MiB: block-relative offsets in the
    decoder need more than 22 bits there (the serial short path once packed them into 22).  Input: random 200-byte runs
    interleaved with short repeats, so every batch mixes long literal runs (general path) with short sequences."""
    import random
    rnd = random.Random(11)
    parts, total = [], 0
    while total < (20 << 20):
        lit = rnd.randbytes(rnd.randrange(150, 260))
        parts.append(lit)
        rep = lit[rnd.randrange(0, 100):][:rnd.randrange(8, 60)]
        for _ in range(rnd.randrange(1, 4)):
            parts.append(rep); parts.append(rnd.randbytes(rnd.randrange(1, 9)))
        total += len(lit) + 4 * len(rep)
    data = b"".join(parts)
    frame = run([lz4_cli, "-c", "-l"], data).stdout
    assert frame[:4] == struct.pack("<I", 0x184C2102)
    first = struct.unpack_from("<I", frame, 4)[0]
    assert first > (4 << 20), f"first legacy block is only {first} B compressed"
    assert run([lz4_cli, "-d", "-c"], frame).stdout == data
    assert run([CLI, "-d", "--all-frames"], frame).stdout == data


def test_cli_linked_blocks_and_dictionary(lz4_cli, gen, tmp_path):
    """-BD (linked blocks) and -D FILE (dictionary): interchangeable with `lz4 -BD` / `lz4 -D` in both directions."""
    dic = gen_bytes(gen, "loglike", 50000, 5)
    dfile = tmp_path / "d.bin"; dfile.write_bytes(dic)
    data = dic[1000:4000] + gen_bytes(gen, "loglike", (5 << 20) + 17, 6)
    fr = run([CLI, "-B4", "-BD"], data).stdout
    assert fr[4] & 0x20 == 0                                        # linked
    assert run([lz4_cli, "-d", "-c"], fr).stdout == data and run([CLI, "-d"], fr).stdout == data
    assert run([CLI, "-d"], run([lz4_cli, "-c", "-B4", "-BD"], data).stdout).stdout == data
    small = dic[2000:2600]
    for payload in (small, data):
        ours = run([CLI, "-B4", "-D", str(dfile)], payload).stdout
        assert run([lz4_cli, "-d", "-c", "-D", str(dfile)], ours).stdout == payload
        assert run([CLI, "-d", "-D", str(dfile)], ours).stdout == payload
        theirs = run([lz4_cli, "-c", "-B4", "-BD", "-D", str(dfile)], payload).stdout
        assert run([CLI, "-d", "-D", str(dfile)], theirs).stdout == payload
    assert len(run([CLI, "-B4", "-D", str(dfile)], small).stdout) < len(run([CLI, "-B4"], small).stdout)


def test_cli_errors_match_the_conduit(lz4_cli):
    frame = run([lz4_cli, "-c"], b"hellohellohellohello" * 1000).stdout
    p = run([CLI, "-d"], frame[:-6], ok=False)
    assert p.returncode != 0 and b"stream ended before EndMark" in p.stderr       # Conduit.hsc:689
    p = run([CLI, "-d"], b"\x01\x02", ok=False)
    assert p.returncode != 0 and b"not enough bytes for header" in p.stderr      # Conduit.hsc:615-616
    p = run([CLI, "-d"], b"\x00" * 64, ok=False)
    assert p.returncode != 0 and b"ERROR_frameType_unknown" in p.stderr
    bad = bytearray(run([lz4_cli, "-c", "-BX"], bytes(range(256)) * 4000).stdout)
    bad[len(bad) // 2] ^= 0x40
    p = run([CLI, "-d"], bytes(bad), ok=False)
    assert p.returncode != 0 and b"lz4frame error: ERROR_" in p.stderr


def test_cli_more_than_32_bit_many_bytes():
    """test/Main.hs:82-86, 105-109 ("5GiB ByteString -- more than 32-bit many Bytes"): streamed through two CLI processes
    in a pipe, nothing of that size is ever resident.  B2_BIG_GIB=0 skips it."""
    gib = float(os.environ.get("B2_BIG_GIB", "5"))
    if gib <= 0:
        pytest.skip("B2_BIG_GIB=0")
    total = int(gib * (1 << 30))
    chunk = bytes([42]) * (64 << 20)
    c = subprocess.Popen([CLI], stdin=subprocess.PIPE, stdout=subprocess.PIPE)
    d = subprocess.Popen([CLI, "-d"], stdin=c.stdout, stdout=subprocess.PIPE)
    c.stdout.close()
    import threading

    def feed():
        left = total
        while left:
            k = min(left, len(chunk))
            c.stdin.write(chunk[:k])
            left -= k
        c.stdin.close()
    t = threading.Thread(target=feed)
    t.start()
    got = 0
    ok = True
    while True:
        buf = d.stdout.read(64 << 20)
        if not buf:
            break
        ok = ok and buf.count(42) == len(buf)
        got += len(buf)
    t.join()
    assert c.wait() == 0 and d.wait() == 0
    assert got == total and ok
