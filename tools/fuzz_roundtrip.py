"""GPU box: randomized round-trip fuzz of the codec against the oracle (test infrastructure).  Inputs are built from random
segments (noise, runs, periodic patterns of random period, copies of earlier data at random distances incl. > 64 KB,
text-like lines), sizes cluster around block / sub-block / unit boundaries; preferences are drawn at random.  For every case:
our frame decodes to the input with the oracle AND with our decoder (with and without the index); every 4th case also goes
through the LZ4F_* layer with random chunk sizes (oracle/conduit_replay.c against the product).
usage: python tools/fuzz_roundtrip.py [cases] [seed]"""
import ctypes as C, importlib, os, random, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import conftest
pkg = importlib.import_module("lz4-frame-conduit_b200"); pkg.load()
sz = C.c_size_t
orc = C.CDLL(os.path.join(ROOT, "oracle", "liboracle_lz4.so"))
orc.orc_frame_decompress.restype = sz
orc.orc_frame_decompress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.POINTER(sz)]
replay = conftest._bind_replay(C.CDLL(conftest.build_replay_b200()))
refp = os.path.join(ROOT, "oracle", "_ref", "libref.so")
ref = conftest._bind_replay(C.CDLL(refp)) if os.path.exists(refp) else None
cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rnd = random.Random(seed); nrg = np.random.default_rng(seed)
WORDS = [b"GET", b"POST", b"error", b"warning", b"user=", b"id=", b" 200 ", b" 404 ", b"/api/v1/items/", b"timeout", b"\n", b" ", b"0123456789"]

def make(n):
    out = bytearray()
    while len(out) < n:
        k = rnd.randrange(7); m = min(n - len(out), rnd.choice([1, 3, 7, 40, 300, 5000, 70000, 300000]))
        if k == 0: out += nrg.integers(0, 256, m, dtype=np.uint8).tobytes()
        elif k == 1: out += bytes([rnd.randrange(256)]) * m
        elif k == 2:
            p = rnd.choice([1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 31, 32, 33, 64, 255, 256, 1000, 4096, 65535, 65536, 65537])
            pat = nrg.integers(0, 256, min(p, m), dtype=np.uint8).tobytes()
            out += (pat * (m // len(pat) + 1))[:m]
        elif k == 3 and out:
            dist = min(len(out), rnd.choice([1, 2, 9, 100, 4000, 65535, 65536, 70000, 200000]))
            st = len(out) - dist
            for i in range(0, m, dist): out += out[st + i % dist: st + i % dist + min(dist, m - i)][:m - i] if dist >= m - i else out[st:st + dist]
            del out[n:]
        elif k == 4: out += nrg.integers(97, 101, m, dtype=np.uint8).tobytes()
        else:
            line = b"".join(rnd.choice(WORDS) for _ in range(rnd.randrange(3, 12))) + b"\n"
            while m > 0: out += line[:m]; m -= len(line); line = line[:rnd.randrange(1, len(line) + 1)] + rnd.choice(WORDS) + b"\n"
    return bytes(out[:n])

def size():
    base = rnd.choice([0, 1, 12, 13, 16384, 65536, 131072, 262144, 1 << 20, 4 << 20])
    return max(0, base * rnd.choice([1, 1, 2, 3]) + rnd.choice([-13, -12, -5, -1, 0, 0, 1, 5, 12, 13, 4097]) + (rnd.randrange(200000) if rnd.random() < 0.3 else 0))

t0 = time.time(); tot = 0
with pkg.Codec(0) as c:
    for i in range(cases):
        n = size(); d = make(n); tot += n
        bid = rnd.choice([4, 5, 6, 7]); sub = min(rnd.choice([16, 17, 18, 20, 22]), 8 + 2 * bid)
        prefs = dict(block_size_id=bid, sub_block_log2=sub, block_checksum=rnd.randrange(2), content_checksum=rnd.randrange(2),
                     content_size=rnd.randrange(2), write_index=rnd.randrange(2), use_index=1)
        c.set_prefs(**prefs)
        fr = c.compress_bytes(d)
        out = C.create_string_buffer(max(1, n)); used = sz(0)
        r = orc.orc_frame_decompress(fr, len(fr), out, n, C.byref(used))
        assert r == n and out.raw[:n] == d, ("oracle rejects / differs", i, n, prefs, r)
        for ui in (1, 0):
            c.set_prefs(use_index=ui)
            got, _ = c.decompress_bytes(fr, max_out=n)
            assert got == d, ("our decoder differs", i, n, prefs, ui)
        if i % 4 == 0:
            cap = n + n // 100 + (70 << 20)
            o = C.create_string_buffer(cap)
            r = replay.ref_conduit_compress(d, n, o, cap, rnd.choice([0, 1000, 32768, 3 << 20]), 0)
            assert r < (1 << 62), ("LZ4F compress failed", i, n)
            fr2 = o.raw[:r]
            used = sz(0)
            r2 = orc.orc_frame_decompress(fr2, len(fr2), out, n, C.byref(used))
            assert r2 == n and out.raw[:n] == d, ("oracle rejects the LZ4F_* frame", i, n)
            got, cons = conftest.ref_conduit_decode(replay, fr2, n + 64, rnd.choice([0, 777, 65536]))
            assert got == d, ("LZ4F decompress differs", i, n)
        if ref is not None and i % 3 == 0:
            # frames the REFERENCE library makes (linked or independent blocks, flushes in odd places): our decoder = the input
            args = (rnd.choice([0, 4, 5, 6, 7]), rnd.randrange(2), rnd.randrange(2), rnd.randrange(2), rnd.randrange(2), rnd.randrange(2))
            chunk = rnd.choice([0, 1, 4000, 65536, 100000, 5 << 20])
            if chunk == 1 and n > 50000: chunk = 4000                 # autoFlush + 1-byte updates = one 9-byte block per input byte
            cap = (10 * n if chunk == 1 else 2 * n) + (1 << 20)
            o = C.create_string_buffer(cap)
            r = ref.ref_frame_compress(d, n, o, cap, *args, chunk)
            assert r < (1 << 62), ("reference compress failed", i, n, args)
            fr3 = o.raw[:r]
            for ui in (1, 0):
                c.set_prefs(use_index=ui)
                got, cons = c.decompress_bytes(fr3, max_out=n)
                assert got == d and cons == len(fr3), ("our decoder differs on a reference-made frame", i, n, args)
            got, cons = conftest.ref_conduit_decode(replay, fr3, n + 64, rnd.choice([0, 777, 65536]))
            assert got == d, ("LZ4F decompress differs on a reference-made frame", i, n, args)
print(f"fuzz ok: {cases} cases, {tot / 1e6:.1f} MB, seed {seed}, {time.time() - t0:.1f} s")
