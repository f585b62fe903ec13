"""GPU box: decode rate of ONE chain (a reference-made frame of linked 64 KB blocks, device-resident) and of an unindexed
frame of few large blocks — the lone-warp cases.  usage: python tools/lone_chain.py"""
import ctypes as C, importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import conftest, torch
pkg = importlib.import_module("lz4-frame-conduit_b200"); pkg.load()
ref = conftest._bind_replay(C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so")))
n = 64 << 20
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so")); gen.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
src = C.create_string_buffer(n); gen.b2gen_loglike(C.addressof(src), n, 0x10C5)
cap = n + (1 << 20); o = C.create_string_buffer(cap)
with pkg.Codec(0) as c:
    for what, args in (("linked 64 KB blocks (1 chain)", (4, 0, 0, 0, 0, 0)), ("independent 4 MB blocks, no index (16 chains)", (7, 1, 0, 0, 0, 0))):
        r = ref.ref_frame_compress(src, n, o, cap, *args, 1 << 20)
        fr = torch.frombuffer(bytearray(o.raw[:r]), dtype=torch.uint8).cuda()
        dst = torch.empty(n, dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
        c.decompress_device(fr.data_ptr(), r, dst.data_ptr(), n)
        t0 = time.perf_counter(); got, _ = c.decompress_device(fr.data_ptr(), r, dst.data_ptr(), n); dt = time.perf_counter() - t0
        assert got == n and bytes(dst[-4096:].cpu().numpy()) == src.raw[n - 4096:n]
        print(f"{what}: {n / dt / 1e9:.3f} GB/s ({dt * 1e3:.0f} ms for 64 MiB)")
