"""GPU box: throughput of the UNMODIFIED conduit's call sequence (oracle/conduit_replay.c: 16 KB LZ4F_compressUpdate calls,
16 KB+ LZ4F_decompress destinations, pageable host buffers) on the product's LZ4F_* symbols, next to the same replay on the
reference library (1 thread).  Also decodes a reference-made (linked 64 KB blocks) frame through the same calls.
usage: python tools/tier1_speed.py [size_mib] [--only-compress]   (B2LZ4_BATCH_MB=.. for a batch-size sweep of the compress legs)"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("B2LZ4_BATCH_MB", "")                      # empty = the library defaults (tests/conftest.py would lower it to 8 for coverage)
import conftest
import importlib
pkg = importlib.import_module("lz4-frame-conduit_b200"); pkg.load()
ours = conftest._bind_replay(C.CDLL(conftest.build_replay_b200()))
refp = os.path.join(ROOT, "oracle", "_ref", "libref.so")
ref = conftest._bind_replay(C.CDLL(refp)) if os.path.exists(refp) else None
only_compress = "--only-compress" in sys.argv
argv = [a for a in sys.argv[1:] if not a.startswith("--")]
n = (int(argv[0]) if argv else 1024) << 20
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so")); gen.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
src = C.create_string_buffer(n); gen.b2gen_loglike(C.addressof(src), n, 0x10C5)
cap = n + n // 100 + (1 << 20)
comp = C.create_string_buffer(cap); back = C.create_string_buffer(n + 64)
def t(f):
    t0 = time.perf_counter(); r = f(); return r, time.perf_counter() - t0
def run(lib, name, size):
    lib.ref_conduit_compress(src, min(size, 64 << 20), comp, cap, 32768, 0)                # warm-up (contexts, pinned staging)
    c, tc = t(lambda: lib.ref_conduit_compress(src, size, comp, cap, 32768, 0))
    assert c < (1 << 62), f"{name}: compress failed with code {(1 << 64) - c}"
    used = C.c_size_t(0)
    d, td = t(lambda: lib.ref_conduit_decompress(comp, c, back, n + 64, 65536, C.byref(used)))
    assert d == size and back.raw[:1 << 20] == src.raw[:1 << 20] and back.raw[size - 4096:size] == src.raw[size - 4096:size], (name, d)
    print(f"{name}: conduit compress {size / tc / 1e9:.2f} GB/s, conduit decompress {size / td / 1e9:.2f} GB/s, frame {c} B (ratio {size / c:.3f})", flush=True)
    return c
run(ours, "b200 LZ4F_* (Tier 1)", n)
# compressYieldImmediately (Conduit.hsc:364-425): a fresh LZ4F_compressBound(len)-sized buffer per upstream ByteString (the replay
# mallocs it like allocaBytes does), here with 32 KB ByteStrings — conduit's default file chunk
ours.ref_conduit_compress_yi(src, min(n, 64 << 20), comp, cap, 32768)
cy, ty = t(lambda: ours.ref_conduit_compress_yi(src, n, comp, cap, 32768))
assert cy < (1 << 62), f"compressYieldImmediately failed with code {(1 << 64) - cy}"
print(f"b200 LZ4F_* (Tier 1): conduit compressYieldImmediately {n / ty / 1e9:.2f} GB/s, frame {cy} B "
      f"(B2LZ4_BATCH_MB={os.environ['B2LZ4_BATCH_MB'] or 'library default'})", flush=True)
if only_compress:
    sys.exit(0)
if ref is not None:
    m = min(n, 256 << 20)
    c = run(ref, "reference library, 1 thread", m)                                          # comp now holds a reference-made frame
    used = C.c_size_t(0)
    d, td = t(lambda: ours.ref_conduit_decompress(comp, c, back, n + 64, 65536, C.byref(used)))
    assert d == m and back.raw[m - 4096:m] == src.raw[m - 4096:m]
    print(f"b200 LZ4F_decompress on the reference-made frame (linked 64 KB blocks, {m >> 20} MiB): {m / td / 1e9:.3f} GB/s", flush=True)

cli = os.path.join(ROOT, "oracle", "_ref", "lz4")
if os.path.exists(cli):
    import subprocess, tempfile
    os.environ.pop("B2LZ4_BATCH_MB", None)                   # library default: batches of large blocks are sized by block count
    m = min(n, 512 << 20)
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as td:
        raw = os.path.join(td, "raw"); open(raw, "wb").write(src.raw[:m])
        for flags, what in ((["-B7", "--no-frame-crc"], "stock CLI frame, 4 MB independent blocks, no content checksum"),
                            ([], "stock CLI default frame (4 MB blocks + content checksum)")):
            fr = subprocess.run([cli, "-c", *flags, raw], capture_output=True, check=True).stdout
            used = C.c_size_t(0)
            for lib, name in ((ours, "b200 LZ4F_decompress"), (ref, "reference library")):
                if lib is None: continue
                # first pass: a context that has never seen such a frame (pins its staging, sizes the dataflow workspaces);
                # second pass: the same process decoding the next such file (staging comes from the pool)
                d, td_ = t(lambda: lib.ref_conduit_decompress(fr, len(fr), back, n + 64, 65536, C.byref(used)))
                assert d == m and back.raw[m - 4096:m] == src.raw[m - 4096:m], (name, d)
                d, tw_ = t(lambda: lib.ref_conduit_decompress(fr, len(fr), back, n + 64, 65536, C.byref(used)))
                assert d == m
                print(f"{name} on a {what} ({m >> 20} MiB): {m / td_ / 1e9:.2f} GB/s first file, {m / tw_ / 1e9:.2f} GB/s next file", flush=True)
