"""GPU box: time b2lz4_compress_host / b2lz4_decompress_host separately (pinned buffers) for the pipeline shape given
by B2LZ4_PIPE_MB / B2LZ4_LANES.  usage: python tools/e2e_sweep.py [size_mib]"""
import ctypes as C, importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
pkg = importlib.import_module("lz4-frame-conduit_b200")
n = (int(sys.argv[1]) if len(sys.argv) > 1 else 1024) << 20
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so")); gen.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
c = pkg.Codec(0, block_size_id=7, sub_block_log2=17, write_index=int(os.environ.get("IDX", "1")))   # the bench shape
h_in = torch.empty(n, dtype=torch.uint8).pin_memory(); gen.b2gen_loglike(h_in.data_ptr(), n, 0x10C5)
cap = pkg.compress_bound(n, c.prefs)
h_c = torch.empty(cap, dtype=torch.uint8).pin_memory(); h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
for _ in range(2):
    cs = c.compress_host_ptr(h_in.data_ptr(), n, h_c.data_ptr(), cap); c.decompress_host_ptr(h_c.data_ptr(), cs, h_b.data_ptr(), n)
assert torch.equal(h_in, h_b)
K = 4
t0 = time.perf_counter()
for _ in range(K): cs = c.compress_host_ptr(h_in.data_ptr(), n, h_c.data_ptr(), cap)
t1 = time.perf_counter(); sc = c.stats()
for _ in range(K): c.decompress_host_ptr(h_c.data_ptr(), cs, h_b.data_ptr(), n)
t2 = time.perf_counter(); sd = c.stats()
print(f"  compress: sum over chunks of find_emit {sc.ms_main:.1f} ms, of whole chunk streams {sc.ms_total:.1f} ms, launches {sc.kernel_launches}; decompress: decode {sd.ms_main:.1f} ms, total {sd.ms_total:.1f} ms")
# plain copies for reference
d = torch.empty(n, dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
t3 = time.perf_counter(); d.copy_(h_in, non_blocking=True); torch.cuda.synchronize(); t4 = time.perf_counter()
h_b.copy_(d, non_blocking=True); torch.cuda.synchronize(); t5 = time.perf_counter()
print(f"PIPE_MB={os.environ.get('B2LZ4_PIPE_MB','auto')} LANES={os.environ.get('B2LZ4_LANES','4')}: compress_host {1e3*(t1-t0)/K:.1f} ms, "
      f"decompress_host {1e3*(t2-t1)/K:.1f} ms, e2e {n*K/(t2-t0)/1e9:.2f} GB/s | H2D {n/(t4-t3)/1e9:.1f} GB/s, D2H {n/(t5-t4)/1e9:.1f} GB/s")
