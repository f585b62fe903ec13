#!/usr/bin/env python
"""Generates tests/golden/* from the UNMODIFIED reference (oracle/_ref, built by `make -C oracle ref`).

Run in the build container only (needs /root/reference through oracle/_ref).  The committed outputs are
what travels: small frames produced by the reference library / the stock lz4 CLI, their payloads, and
a manifest with sizes and XXH32 values computed by the reference's own xxhash.c.
"""
import ctypes as C
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref")

ref = C.CDLL(os.path.join(REF, "libref.so"))
sz = C.c_size_t
ref.ref_conduit_compress.restype = sz
ref.ref_conduit_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz, sz, sz]
ref.ref_frame_compress.restype = sz
ref.ref_frame_compress.argtypes = [C.c_char_p, sz, C.c_char_p, sz] + [C.c_int] * 6 + [sz]
ref.ref_xxh32.restype = C.c_uint
ref.ref_xxh32.argtypes = [C.c_char_p, sz, C.c_uint]
ref.ref_frame_compress_dict.restype = sz
ref.ref_frame_compress_dict.argtypes = [C.c_char_p, sz, C.c_char_p, sz, C.c_char_p, sz, C.c_int, C.c_int, C.c_int, C.c_uint, sz]


def payloads():
    import random
    rnd = random.Random(7)
    p = {}
    p["empty"] = b""
    p["one"] = b"x"
    p["hello"] = b"hellohellohellohello"                                   # test/Main.hs:61
    p["twelve"] = b"abcabcabcabc"                                          # < 13 bytes: literal-only (lz4.c:859)
    p["thirteen"] = b"abcabcabcabca"
    p["numbers"] = ("BEGIN " + " ".join(str(i) for i in range(1, 20001)) + " END").encode()   # test/Main.hs:66 (scaled)
    p["hellos"] = b"BEGIN" + b"hello" * 20000 + b"END"                      # test/Main.hs:71 (scaled)
    p["bytes42"] = bytes([42]) * 100000                                    # test/Main.hs:76
    p["random70k"] = bytes(rnd.getrandbits(8) for _ in range(70000))       # stored blocks
    p["b65535"] = p["numbers"][:65535]
    p["b65536"] = p["numbers"][:65536]
    p["b65537"] = p["numbers"][:65537]
    dg = subprocess.run([os.path.join(REF, "datagen"), "-g200000", "-s1"], capture_output=True, check=True).stdout
    assert len(dg) == 200000
    p["datagen200k"] = dg                                                  # the reference's own generator
    mixed = dg[:30000] + p["random70k"][:66000] + bytes(40000) + dg[30000:60000]
    p["mixed166k"] = mixed
    return p


def conduit_frame(data):
    cap = len(data) + len(data) // 100 + 65536 * 2
    out = C.create_string_buffer(cap)
    r = ref.ref_conduit_compress(data, len(data), out, cap, 0, 0)
    assert r < (1 << 62), r
    return out.raw[:r]


def lib_frame(data, bsid, indep, cc, bc, cs, autoflush=1, chunk=0):
    cap = len(data) + len(data) // 100 + 65536 * 2 + (len(data) // 1024 + 1) * 8
    out = C.create_string_buffer(cap)
    r = ref.ref_frame_compress(data, len(data), out, cap, bsid, indep, cc, bc, cs, autoflush, chunk)
    assert r < (1 << 62), r
    return out.raw[:r]


def cli_frame(data, *args):
    r = subprocess.run([os.path.join(REF, "lz4"), "-c", *args], input=data, capture_output=True)
    if r.returncode != 0:      # the vendored v1.9.2-dev CLI fails on 65535 bytes with -B4 -BX (its own bound check)
        print("skipped: lz4 -c", *args, "on", len(data), "bytes:", r.stderr.decode().strip().splitlines()[-1])
        return None
    return r.stdout


def main():
    man = {}
    ps = payloads()
    for name, data in ps.items():
        boundary = name.startswith("b6553")       # boundary sizes: payload = numbers.raw[:n], fewer frame kinds
        if not boundary:
            with open(os.path.join(HERE, name + ".raw"), "wb") as f:
                f.write(data)
        frames = {"conduit": conduit_frame(data),                      # reference conduit defaults: linked 64 KB
                  "cli": cli_frame(data),                               # lz4 -c: independent 4 MB + content checksum
                  "cli_B4_BX": cli_frame(data, "-B4", "-BX"),           # 64 KB independent + block + content checksums
                  "cli_B4_BD": cli_frame(data, "-B4", "-BD", "--no-frame-crc"),   # linked 64 KB (ext-dict parse)
                  "lib_B5_cs": lib_frame(data, 5, 1, 1, 0, 1, 1, 256 * 1024),      # 256 KB independent + contentSize
                  "lib_B4_flush": lib_frame(data, 4, 1, 0, 1, 0, 1, 40000)}        # autoFlush with 40000-byte updates:
        #                                                                 short blocks in the middle of the frame
        if boundary:
            frames = {k: frames[k] for k in ("conduit", "cli_B4_BX", "cli_B4_BD")}
        ent = {"size": len(data), "raw": "numbers.raw" if boundary else name + ".raw", "xxh32": ref.ref_xxh32(data, len(data), 0), "frames": {}}
        for kind, fr in frames.items():
            if fr is None:
                continue
            fn = f"{name}.{kind}.lz4"
            with open(os.path.join(HERE, fn), "wb") as f:
                f.write(fr)
            ent["frames"][kind] = {"file": fn, "size": len(fr), "xxh32": ref.ref_xxh32(fr, len(fr), 0),
                                   "head": fr[:7].hex()}
        man[name] = ent
    # ---- dictionary frames (LZ4F_compressBegin_usingCDict, lz4frame.c:594-702; lz4/tests/frametest.c:520-662): linked blocks,
    # the dictionary = 40 KB of the reference's generator with another seed; with and without a dictID in the header
    dic = subprocess.run([os.path.join(REF, "datagen"), "-g40000", "-s1"], capture_output=True, check=True).stdout[:40000]
    with open(os.path.join(HERE, "dict40k.raw"), "wb") as f:
        f.write(dic)
    dman = {"file": "dict40k.raw", "size": len(dic), "frames": {}}
    for name, did in (("hello", 0), ("numbers", 0x00C0FFEE), ("datagen200k", 0), ("mixed166k", 7)):
        data = ps[name]
        cap = len(data) + len(data) // 50 + 70000
        out = C.create_string_buffer(cap)
        r = ref.ref_frame_compress_dict(data, len(data), out, cap, dic, len(dic), 4, 0, 1, did, 16384)
        assert r < (1 << 62), r
        fn = f"{name}.lib_dict.lz4"
        with open(os.path.join(HERE, fn), "wb") as f:
            f.write(out.raw[:r])
        dman["frames"][name] = {"file": fn, "size": r, "dict_id": did, "head": out.raw[:11].hex()}
    man["_dict"] = dman
    man["_xxh32_kat"] = {"": ref.ref_xxh32(b"", 0, 0), "abc": ref.ref_xxh32(b"abc", 3, 0),
                         "hello*4": ref.ref_xxh32(b"hellohellohellohello", 20, 0)}
    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump(man, f, indent=1, sort_keys=True)
    tot = sum(os.path.getsize(os.path.join(HERE, x)) for x in os.listdir(HERE))
    print("wrote", len(man) - 1, "payloads,", tot, "bytes total")


if __name__ == "__main__":
    sys.exit(main())
