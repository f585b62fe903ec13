import os, sys, struct, importlib, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["B2LZ4_PIPE_MB"] = "8"
pkg = importlib.import_module("lz4-frame-conduit_b200")
from conftest import gen_bytes
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so"))
for n in ("b2gen_loglike", "b2gen_random", "b2gen_textlike"):
    getattr(gen, n).argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]; getattr(gen, n).restype = None
data = gen_bytes(gen, "loglike", (150 << 20) + 12345) + gen_bytes(gen, "random", 9 << 20, 7) + bytes(13 << 20)
kw = dict(write_index=1, content_checksum=1, block_size_id=6)
with pkg.Codec(0, **kw) as one, pkg.Codec(0, **kw) as many:
    many.set_devices([0, 0, 0])
    f1 = one.compress_bytes(data); f2 = many.compress_bytes(data); f3 = many.compress_bytes(data); f4 = one.compress_bytes(data)
    print(len(f1), len(f2), f1 == f4, f2 == f3)
    diffs = [i for i in range(0, min(len(f1), len(f2))) if f1[i] != f2[i]][:20] if f1 != f2 else []
    print("diffs", len(diffs), diffs[:20])
    # layout
    pb = struct.unpack("<I", f1[-8:-4])[0]; ix = len(f1) - 16 - pb
    print("index frame at", ix, "payload", pb, "endmark at", ix - 8, "checksum", f1[ix - 4:ix].hex(), f2[ix - 4:ix].hex())
    h = struct.unpack("<6IQ", f1[ix + 8: ix + 8 + 32]); print("head", h)
    nb, nc = h[3], h[4]
    cb = ix + 8 + 32; cc = cb + nb * 16
    for d in diffs[:6]:
        if d >= cc: print(d, "in chain entry", (d - cc) // 8, "byte", (d - cc) % 8)
        elif d >= cb: print(d, "in block entry", (d - cb) // 16, "byte", (d - cb) % 16)
        else: print(d, "before index")
    for name, f in (("one", f1), ("many", f2)):
        print(name, [struct.unpack("<4I", f[cb + 16 * k: cb + 16 * k + 16]) for k in (0, 1, 2, 3, 4)])
    # real header words
    off = 7
    words = []
    for k in range(10):
        w = struct.unpack("<I", f1[off:off + 4])[0]; words.append(w); off += 4 + (w & 0x7fffffff)
    print("real", words)
