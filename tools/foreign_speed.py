"""GPU box: device-resident decompression of frames this codec did NOT write (no index): stock-CLI frames (4 MB independent
blocks, 64 KB linked blocks), a reference-conduit frame (linked 64 KB, Conduit.hsc:248-263) and this codec's own plain frames.
These go through the dataflow kernels (csrc/decode_df.cu); B2LZ4_DF=0 selects the one-warp-per-block path for comparison.
usage: python tools/foreign_speed.py [size_mib] [workload]"""
import ctypes as C, importlib, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import conftest
pkg = importlib.import_module("lz4-frame-conduit_b200")
n = (int(sys.argv[1]) if len(sys.argv) > 1 else 1024) << 20
workload = sys.argv[2] if len(sys.argv) > 2 else "loglike"
gen = C.CDLL(os.path.join(ROOT, "tools", "libb2gen.so"))
host = torch.empty(n, dtype=torch.uint8).pin_memory()
if workload == "zeros":
    host.zero_()
else:
    f = getattr(gen, "b2gen_" + workload); f.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]; f.restype = None
    f(host.data_ptr(), n, 0x10C5)
raw = bytes(memoryview(host.numpy()))
cli = os.path.join(ROOT, "oracle", "_ref", "lz4")
frames = {}
with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as td:
    p = os.path.join(td, "raw"); open(p, "wb").write(raw)
    for name, flags in (("cli_B7_indep", ["-B7", "--no-frame-crc"]), ("cli_B4_linked", ["-B4", "-BD", "--no-frame-crc"]),
                        ("cli_B4_indep", ["-B4", "--no-frame-crc"]), ("cli_default_crc", [])):
        frames[name] = subprocess.run([cli, "-c", *flags, p], capture_output=True, check=True).stdout
refp = os.path.join(ROOT, "oracle", "_ref", "libref.so")
if os.path.exists(refp):
    ref = conftest._bind_replay(C.CDLL(refp))
    cap = n + n // 100 + (1 << 20)
    comp = C.create_string_buffer(cap)
    c = ref.ref_conduit_compress(raw, n, comp, cap, 32768, 0)
    frames["ref_conduit_linked"] = comp.raw[:c]
d_in = host.cuda()
d_out = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
res = {"size_mib": n >> 20, "workload": workload, "df": os.environ.get("B2LZ4_DF", "1")}
with pkg.Codec(0, use_index=0) as c:
    for bid, tag in ((5, "own_plain_B5"), (7, "own_plain_B7")):
        c.set_prefs(block_size_id=bid, sub_block_log2=18 if bid == 5 else 17, write_index=0)
        cap = pkg.compress_bound(n, c.prefs)
        d_c = torch.empty(cap, dtype=torch.uint8, device="cuda")
        sz = c.compress_device(d_in.data_ptr(), n, d_c.data_ptr(), cap)
        frames[tag] = bytes(memoryview(d_c[:sz].cpu().numpy()))
        del d_c
    for name, fr in frames.items():
        d_f = torch.frombuffer(bytearray(fr), dtype=torch.uint8).cuda()
        best = 1e9; main = 0
        for it in range(5):
            d_out.zero_(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            r, used = c.decompress_device(d_f.data_ptr(), len(fr), d_out.data_ptr(), n)
            dt = time.perf_counter() - t0
            st = c.stats()
            if st.ms_total * 1e-3 < best: best, main = st.ms_total * 1e-3, st.ms_main
            assert r == n and used == len(fr), (name, r, used)
            if it == 0:
                assert torch.equal(d_out[:n], d_in), name + ": output differs"
        res[name] = {"GBps": round(n / best / 1e9, 2), "ms_total": round(best * 1e3, 3), "ms_decode": round(main, 3),
                     "ratio": round(n / len(fr), 3), "blocks": st.n_blocks, "pieces": st.n_chains, "retries": st.retries, "wall_ms_last": round(dt * 1e3, 2)}
        if os.environ.get("B2_DF_PROF"):
            pr = (C.c_ulonglong * 12)()
            c.lib.b2lz4_debug_df_prof.argtypes = [C.c_void_p, C.c_void_p]; c.lib.b2lz4_debug_df_prof.restype = C.c_size_t
            c.lib.b2lz4_debug_df_prof(c.h, pr)
            if pr[3]:
                res[name]["prof"] = {"tickets": pr[3], "cyc_flow_per_ticket": pr[0] / pr[3], "cyc_decode_per_ticket": pr[1] / pr[3],
                                     "cyc_retire_per_ticket": pr[2] / pr[3], "rounds_per_ticket": pr[4] / pr[3], "idle_rounds_per_ticket": pr[5] / pr[3], "cyc_parse": pr[6] / pr[3], "cyc_literals": pr[7] / pr[3], "cyc_chase": pr[8] / pr[3], "cyc_rounds": pr[9] / pr[3]}
        print(name, res[name], flush=True)
print(json.dumps(res))
