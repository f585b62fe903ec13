"""Compressed size of the datagen (text-like) and log-like shapes against the reference's default preferences, for the
library B2LZ4_LIB selects (tuning builds).  Needs a GPU and oracle/_ref (test infrastructure: the reference is the checker)."""
import ctypes as C, importlib, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("lz4-frame-conduit_b200")

ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so")) if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref.so")) else None
out = []
with pkg.Codec(0) as c:
    c.set_prefs(block_size_id=7, block_checksum=0, content_checksum=0, content_size=0, sub_block_log2=18, write_index=0)
    t = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "datagen"), "-g16777216", "-s1"], capture_output=True, check=True).stdout
    for name, d in (("datagen", t),):
        ours = len(c.compress_bytes(d))
        if ref is not None:
            ref.ref_conduit_compress.restype = C.c_size_t
            ref.ref_conduit_compress.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_size_t, C.c_size_t]
            cap = len(d) + len(d) // 100 + 100000
            o = C.create_string_buffer(cap)
            r = ref.ref_conduit_compress(d, len(d), o, cap, 0, 0)
            out.append(f"{name} ours {ours} ref {r} ({100.0 * (ours / r - 1):+.2f} %)")
        else:
            out.append(f"{name} ours {ours}")
print("  ".join(out))
