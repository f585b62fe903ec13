"""BASELINE.json config 3 shape: one LZ4 frame of more than 2^32 bytes streamed end to end through the host-buffer C ABI
(pinned host in -> frame in pinned host memory -> pinned host out), its blocks sharded over every GPU of the box by
block index (b2lz4_set_devices), reassembled in order on the host.   usage: python tools/config3.py [GiB=5] [ngpus=all]
Prints one JSON line (wall clock, copies included)."""
import ctypes as C, importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("lz4-frame-conduit_b200")
gib = float(sys.argv[1]) if len(sys.argv) > 1 else 5.0
ngpu = int(sys.argv[2]) if len(sys.argv) > 2 else torch.cuda.device_count()
n = int(gib * (1 << 30))
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so")); gen.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
c = pkg.Codec(0, content_checksum=int(os.environ.get("SUM", "0")))     # library defaults: plain frame, 256 KB blocks
c.set_devices(list(range(ngpu)))
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
for i in range(0, n, 1 << 30):                                   # per-GiB reseeding
    gen.b2gen_loglike(h_in.data_ptr() + i, min(1 << 30, n - i), 0x10C5 + (i >> 30))
cap = pkg.compress_bound(n, c.prefs)
h_c = torch.empty(cap, dtype=torch.uint8).pin_memory(); h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
cs = c.compress_host_ptr(h_in.data_ptr(), n, h_c.data_ptr(), cap); c.decompress_host_ptr(h_c.data_ptr(), cs, h_b.data_ptr(), n)   # warm-up: lanes, workspaces
K = 2
t0 = time.perf_counter()
for _ in range(K): cs = c.compress_host_ptr(h_in.data_ptr(), n, h_c.data_ptr(), cap)
t1 = time.perf_counter()
for _ in range(K): r, used = c.decompress_host_ptr(h_c.data_ptr(), cs, h_b.data_ptr(), n)
t2 = time.perf_counter()
ok = (r == n and used == cs)
step = 1 << 28
for i in range(0, n, step):
    ok = ok and torch.equal(h_in[i:i + step], h_b[i:i + step])
bs = (1 << (8 + 2 * c.prefs.block_size_id)) >> 10
print(json.dumps({"workload": f"config3 shape: one frame of {gib} GiB log-like (per-GiB reseeded), {bs} KB independent blocks, host in -> host out",
                  "n_gpus": ngpu, "bytes": n, "frame_bytes": cs, "ratio": n / cs, "bit_exact": bool(ok),
                  "compress_host_GBps": n * K / (t1 - t0) / 1e9, "decompress_host_GBps": n * K / (t2 - t1) / 1e9,
                  "roundtrip_GBps": n * K / (t2 - t0) / 1e9, "content_checksum": int(os.environ.get("SUM", "0"))}))
