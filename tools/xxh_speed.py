"""GPU box: throughput of the streamed content checksum (b2lz4_xxh32_device, 1 GiB)."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, ctypes as C
pkg = importlib.import_module("lz4-frame-conduit_b200")
c = pkg.Codec(0)
n = 1 << 30
torch.manual_seed(1)
d = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
h = C.c_uint32(0)
c.lib.b2lz4_xxh32_device(c.h, C.c_void_p(d.data_ptr()), C.c_size_t(n), C.byref(h)); torch.cuda.synchronize()
t0 = time.perf_counter(); c.lib.b2lz4_xxh32_device(c.h, C.c_void_p(d.data_ptr()), C.c_size_t(n), C.byref(h)); t1 = time.perf_counter()
print(f"xxh32 stream: {n / (t1 - t0) / 1e9:.2f} GB/s  digest {h.value:08x}")
