#!/usr/bin/env python
"""bench.py — LZ4 frame compress + decompress throughput on B200 (BASELINE.json metric).

A "step" is one pass of the hot path over one batch: compress the synthetic stream into ONE LZ4 frame
and decompress that frame again.  `value` = uncompressed bytes through the round trip per second
(U / (t_compress + t_decompress)), device-resident; `e2e` = the same through the host-buffer C-ABI
calls (pinned host buffers, H2D + D2H inside the timed region).  compress / decompress GB/s are also
reported separately, as are the rooflines of the two dominant kernels.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--size-mib M]
N > 1: launched by torchrun, one rank per GPU, every rank its own stream (weak scaling, no collective on
the data path: blocks are independent).
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "lz4f_roundtrip_uncompressed_GBps"
UNIT = "GB/s"
WORKLOAD = "config2: 1 GiB synthetic log-like stream (tools/b2gen.c, splitmix64 seed 0x10C5), 4 MB independent blocks, compress + decompress"
# the other BASELINE.json shapes (parity-test cases; measured with --workload for profiles/, never the default line)
WORKLOADS = {
    "loglike": WORKLOAD,
    "textlike": "config1 shape: text-like stream (tools/b2gen.c b2gen_textlike, datagen-style 50 % match probability), compress + decompress",
    "zeros": "config5 shape: all-zero stream (one long overlapping match per block), compress + decompress",
    "random": "config5 shape: incompressible random stream (every block stored), compress + decompress",
}


def fill(gen, ptr, n, workload, seed):
    if workload == "loglike":
        gen.b2gen_loglike(ptr, n, seed)
    elif workload == "textlike":
        gen.b2gen_textlike(ptr, n, seed)
    elif workload == "random":
        gen.b2gen_random(ptr, n, seed)
    elif workload == "zeros":
        C.memset(ptr, 0, n)
    else:
        raise SystemExit(f"unknown workload {workload}")


def load_gen():
    path = os.path.join(ROOT, "tools", "libb2gen.so")
    if not os.path.exists(path):
        subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", path, os.path.join(ROOT, "tools", "b2gen.c")], check=True)
    lib = C.CDLL(path)
    lib.b2gen_loglike.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
    lib.b2gen_loglike.restype = None
    for f in (lib.b2gen_textlike, lib.b2gen_random):
        f.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64]
        f.restype = None
    return lib


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx = max(mx, float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ reference arm
def ref_lib():
    path = os.path.join(ROOT, "oracle", "_ref", "libref.so")
    if not os.path.exists(path):
        if os.path.isdir("/root/reference/lz4/lib"):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, capture_output=True)
        else:
            return None
    lib = C.CDLL(path)
    sz = C.c_size_t
    lib.ref_conduit_compress.restype = sz
    lib.ref_conduit_compress.argtypes = [C.c_void_p, sz, C.c_void_p, sz, sz, sz]
    lib.ref_conduit_decompress.restype = sz
    lib.ref_conduit_decompress.argtypes = [C.c_void_p, sz, C.c_void_p, sz, sz, C.POINTER(sz)]
    return lib


def ref_roundtrip(lib, src_ptr, n, threads):
    """The reference's CPU conduit (oracle/_ref: unmodified lz4 library + replay of Conduit.hsc's call
    sequence) over `threads` independent slices, one OS thread each (ctypes releases the GIL).
    Returns (t_compress, t_decompress, compressed_bytes)."""
    import numpy as np
    per = (n + threads - 1) // threads
    slices = [(i * per, min(per, n - i * per)) for i in range(threads) if i * per < n]
    comp = [np.empty(ln + ln // 100 + 200000, np.uint8) for _, ln in slices]
    back = [np.empty(ln + 64, np.uint8) for _, ln in slices]
    csz = [0] * len(slices)
    ok = [True] * len(slices)

    def run(fn):
        ts = [threading.Thread(target=fn, args=(i,)) for i in range(len(slices))]
        t0 = time.perf_counter()
        for t in ts: t.start()
        for t in ts: t.join()
        return time.perf_counter() - t0

    def do_c(i):
        off, ln = slices[i]
        csz[i] = lib.ref_conduit_compress(src_ptr + off, ln, comp[i].ctypes.data, comp[i].size, 32768, 0)

    def do_d(i):
        off, ln = slices[i]
        consumed = C.c_size_t(0)
        r = lib.ref_conduit_decompress(comp[i].ctypes.data, csz[i], back[i].ctypes.data, back[i].size, 32768, C.byref(consumed))
        ok[i] = (r == ln)

    tc = run(do_c)
    td = run(do_d)
    assert all(ok) and all(c < (1 << 62) for c in csz), "reference round trip failed"
    return tc, td, sum(csz)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    lib = ref_lib()
    import numpy as np
    gen = load_gen()
    n = args.size_mib << 20
    threads = os.cpu_count() or 1
    sample = n                                                    # the whole stream of the workload (1 GiB: well under a second per step on all cores)
    buf = np.empty(sample, np.uint8)
    gen.b2gen_loglike(buf.ctypes.data, sample, 0x10C5)
    if lib is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref.so not built and /root/reference absent"}))
        return 0
    for _ in range(args.warmup):
        ref_roundtrip(lib, buf.ctypes.data, min(sample, 64 << 20), threads)
    t0 = time.perf_counter()
    tcs = tds = 0.0
    csize = 0
    for _ in range(args.steps):
        tc, td, csize = ref_roundtrip(lib, buf.ctypes.data, sample, threads)
        tcs += tc; tds += td
    wall = time.perf_counter() - t0
    val = sample * args.steps / (tcs + tds) / 1e9
    # the reference conduit is single-threaded (Conduit.hsc has no parallelism): its real figure, on a bounded sample
    one = min(sample, 256 << 20)
    tc1, td1, _ = ref_roundtrip(lib, buf.ctypes.data, one, 1)
    line = {"metric": METRIC, "value": val, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * (tcs + tds) / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "bytes_per_gpu": n},
            "reference_setup": f"whole {sample >> 20} MiB stream per step cut into {threads} independent slices, one reference conduit (linked 64 KB blocks) per host thread",
            "one_thread": {"value": one / (tc1 + td1) / 1e9, "unit": UNIT, "compress_GBps": one / tc1 / 1e9, "decompress_GBps": one / td1 / 1e9,
                           "sample": f"first {one >> 20} MiB, 1 thread: what the reference conduit really is"},
            "compress_GBps": sample * args.steps / tcs / 1e9, "decompress_GBps": sample * args.steps / tds / 1e9,
            "ratio": sample / csize,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "reference",
                             "sample": f"whole {sample >> 20} MiB log-like stream, reference conduit replay (linked 64 KB blocks), {threads} threads on independent slices"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": wall}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------ our arm
def max_over_ranks(values, device="cpu"):
    """Element-wise MAX of per-rank timings over the process group (identity without one): the job is as slow as its
    slowest rank.  Works on NCCL (device tensors) and gloo (CPU tensors, used by the world_size-2 CPU test)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def stream_seed(rank):
    """Every rank compresses its own stream (weak scaling): seed of rank r's synthetic input."""
    return 0x10C5 + rank


def link_bandwidth(torch, n, barrier=None):
    """Pinned host <-> device copy bandwidth of this GPU's host link (the end-to-end roofline), GB/s.  With several
    ranks every rank copies at the same time (barrier in front of each direction) and the SLOWEST rank's figure is
    returned: the ranks share the host's PCIe fabric and memory, and the job is timed as max over ranks.
    `duplex` = per-direction rate with both directions busy at once."""
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
    d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()
    out = {}
    for name in ("h2d", "d2h", "duplex"):
        d.copy_(h, non_blocking=True); h2.copy_(d2, non_blocking=True); torch.cuda.synchronize()
        if barrier:
            barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if name in ("d2h", "duplex"):
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                h2.copy_(d2, non_blocking=True)
        if name in ("h2d", "duplex"):
            d.copy_(h, non_blocking=True)
        torch.cuda.current_stream().wait_stream(side)
        e1.record(); torch.cuda.synchronize()
        ms = max_over_ranks([e0.elapsed_time(e1)], "cuda")[0]
        out[name] = n / (ms * 1e-3) / 1e9
    return out


def foreign_leg(torch, pkg, codec, lib, host_in, d_in, n):
    """A frame this codec did not write: the reference library's own LZ4F output with the stock CLI's shape (4 MB independent
    blocks, lz4io.c:159-175; no checksums), made on the host from the same stream.  Decompressed device-resident and end to
    end.  No index: it goes through the dataflow kernels (csrc/decode_df.cu)."""
    import numpy as np
    if lib is None:
        return None
    lib.ref_frame_compress.restype = C.c_size_t
    lib.ref_frame_compress.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t] + [C.c_int] * 6 + [C.c_size_t]
    cap = n + n // 200 + (1 << 20)
    host_f = torch.empty(cap, dtype=torch.uint8).pin_memory()
    t0 = time.perf_counter()
    fs = lib.ref_frame_compress(host_in.data_ptr(), n, host_f.data_ptr(), cap, 7, 1, 0, 0, 0, 1, 4 << 20)
    t_make = time.perf_counter() - t0
    if fs > (1 << 62):
        return {"error": "reference library failed to make the frame"}
    d_f = host_f[:fs].cuda()
    d_o = torch.empty(n, dtype=torch.uint8, device="cuda")
    best = None
    for it in range(4):
        r, used = codec.decompress_device(d_f.data_ptr(), fs, d_o.data_ptr(), n)
        st = codec.stats()
        assert r == n and used == fs
        if it == 0:
            assert torch.equal(d_o, d_in), "foreign frame: output differs"
        elif best is None or st.ms_total < best.ms_total:
            best = st
    host_o = torch.empty(n, dtype=torch.uint8).pin_memory()
    codec.decompress_host_ptr(host_f.data_ptr(), fs, host_o.data_ptr(), n)
    t0 = time.perf_counter()
    r, used = codec.decompress_host_ptr(host_f.data_ptr(), fs, host_o.data_ptr(), n)
    t_host = time.perf_counter() - t0
    ok = r == n and torch.equal(host_o, host_in)
    return {"frame": "reference library LZ4F output, 4 MB independent blocks, no checksums, no index (the stock CLI's shape)",
            "frame_bytes": fs, "ratio": n / fs, "made_on_host_s": t_make,
            "decompress_GBps": n / (best.ms_total * 1e-3) / 1e9, "decode_kernels_ms": best.ms_main, "ms_total": best.ms_total,
            "tickets": best.n_chains, "blocks": best.n_blocks,
            "e2e_decompress_GBps": n / t_host / 1e9, "bit_exact": bool(ok)}


def sharded_leg(torch, pkg, gen, n_dev, seg_mib=1024):
    """BASELINE.json configs 3 and 5 through ONE context whose host-buffer calls shard the blocks of ONE frame over n_dev
    GPUs by block index (b2lz4_set_devices; no collective: blocks are independent) and reassemble them in order on the
    host.  The stream is never resident: it is produced, compressed, decompressed and compared one pinned segment at a time
    (b2lz4_compress_blocks_host / b2lz4_decompress_blocks_host; the frame header and EndMark are a few host-written bytes).
    `frame_xxh32` is the XXH32 of the compressed frame: it must not depend on the number of devices."""
    import numpy as np
    try:
        import xxhash
    except Exception:
        xxhash = None
    lib = pkg.load()
    seg = seg_mib << 20
    out = {"n_devices": n_dev, "segment_mib": seg_mib,
           "api": "b2lz4_set_devices + b2lz4_compress_blocks_host / b2lz4_decompress_blocks_host, pinned host segments"}
    codec = pkg.Codec(0, write_index=0, use_index=0)          # library defaults (b2lz4_default_prefs): plain frame, 256 KB blocks
    codec.set_devices(list(range(n_dev)))
    cap = pkg.compress_bound(seg, codec.prefs)
    h_in = torch.empty(seg, dtype=torch.uint8).pin_memory()
    h_c = [torch.empty(cap, dtype=torch.uint8).pin_memory() for _ in range(2)]
    h_b = torch.empty(seg, dtype=torch.uint8).pin_memory()
    info = pkg.FrameInfo()
    hdr = pkg.frame_header_write(codec.prefs, 0)
    lib.b2lz4_frame_header_parse(hdr, len(hdr), C.byref(info))
    threads = min(16, os.cpu_count() or 1)

    def fill_seg(kind, i):
        if kind == "zeros":
            if i == 0:
                h_in.zero_()
            return
        # always 16 slices with their own seeds, whatever the number of host cores: the stream (and with it `frame_xxh32`)
        # must be the same on every box
        slices = 16
        per = seg // slices
        fn = gen.b2gen_loglike if kind == "loglike" else gen.b2gen_random
        def work(k0):
            for k in range(k0, slices, threads):
                fn(h_in.data_ptr() + k * per, per if k + 1 < slices else seg - k * per, (0x10C5 + i) * 64 + k)
        ts = [threading.Thread(target=work, args=(k0,)) for k0 in range(threads)]
        for t in ts: t.start()
        for t in ts: t.join()

    def stream(kind, gib):
        nseg = (gib << 30) // seg
        x = xxhash.xxh32() if xxhash else None
        if x: x.update(hdr)
        tc = td = 0.0; comp_total = len(hdr); ok = True
        for i in range(nseg):
            fill_seg(kind, i)
            hc = h_c[i & 1]
            flags = 1 if i == 0 else 0
            t0 = time.perf_counter()
            c = lib.b2lz4_compress_blocks_host(codec.h, h_in.data_ptr(), seg, hc.data_ptr(), cap, flags)
            tc += time.perf_counter() - t0
            if lib.b2lz4_is_error(c):
                return {"error": lib.b2lz4_error_name(c).decode() + " (compress) " + lib.b2lz4_last_cuda_error(codec.h).decode()}
            if x: x.update(memoryview(hc.numpy())[:c])
            comp_total += c
            t0 = time.perf_counter()
            r = lib.b2lz4_decompress_blocks_host(codec.h, C.byref(info), hc.data_ptr(), c, h_b.data_ptr(), seg, flags)
            td += time.perf_counter() - t0
            if lib.b2lz4_is_error(r):
                return {"error": lib.b2lz4_error_name(r).decode() + " (decompress) " + lib.b2lz4_last_cuda_error(codec.h).decode()}
            ok = ok and r == seg and torch.equal(h_b, h_in)
        end = C.create_string_buffer(8)
        e = lib.b2lz4_frame_end(codec.h, end, 8)
        if x: x.update(end.raw[:e])
        comp_total += e
        U = nseg * seg
        return {"bytes": U, "frame_bytes": comp_total, "ratio": U / comp_total, "bit_exact": bool(ok),
                "compress_GBps": U / tc / 1e9, "decompress_GBps": U / td / 1e9, "roundtrip_GBps": U / (tc + td) / 1e9,
                "frame_xxh32": (x.intdigest() if x else None)}

    stream("loglike", 1)                                              # warm-up: lanes on every device, workspaces, pinned staging
    out["config3"] = dict(stream("loglike", 16), workload="ONE frame of 16 GiB log-like (> 2^32 bytes), 256 KB independent blocks (library default), host in -> host out")
    out["config5_random"] = dict(stream("random", 4), workload="4 GiB incompressible random: every block stored (lz4frame.c:749-752)")
    out["config5_zeros"] = dict(stream("zeros", 4), workload="4 GiB zeros: one long overlapping match per block (lz4.c:425-456)")
    codec.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--size-mib", type=int, default=1024)
    ap.add_argument("--sub-log2", type=int, default=17)
    ap.add_argument("--no-index", action="store_true", help="plain frames (drop-in default) instead of seekable ones")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-foreign", action="store_true", help="no foreign-frame (stock CLI shape) decompression leg")
    ap.add_argument("--skip-sharded", action="store_true", help="no in-process multi-GPU leg (configs 3 and 5)")
    ap.add_argument("--linked", action="store_true", help="linked blocks (blockMode = linked, SURVEY 8 f4): not the BASELINE configuration")
    ap.add_argument("--content-checksum", action="store_true", help="frames with the content XXH32 (the stock CLI's default): not the BASELINE configuration")
    ap.add_argument("--block-checksum", action="store_true", help="frames with per-block XXH32 (lz4 -BX): not the BASELINE configuration")
    ap.add_argument("--workload", default="loglike", choices=sorted(WORKLOADS))
    ap.add_argument("--block-id", type=int, default=7, help="frame block size id 4..7 = 64 KB .. 4 MB")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        print("bench.py: no CUDA device — this benchmark has no CPU path", file=sys.stderr)
        return 2
    torch.cuda.set_device(local)
    side = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        side = dist.new_group(backend="gloo")       # host-side barrier for the leg where only rank 0 works (an NCCL barrier spins on the GPUs)

    pkg = importlib.import_module("lz4-frame-conduit_b200")
    gen = load_gen()
    n = args.size_mib << 20
    index = 0 if args.no_index else 1
    codec = pkg.Codec(local, block_size_id=args.block_id, sub_block_log2=min(args.sub_log2, 8 + 2 * args.block_id),
                      write_index=index, use_index=1, block_checksum=1 if args.block_checksum else 0,
                      block_linked=1 if args.linked else 0, content_checksum=1 if args.content_checksum else 0)
    workload = WORKLOADS[args.workload]
    block_size = 1 << (8 + 2 * args.block_id)

    # synthetic input: every rank its own stream (weak scaling), generated on the host, moved to HBM once
    host_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    fill(gen, host_in.data_ptr(), n, args.workload, stream_seed(rank))
    d_in = host_in.cuda()
    cap = pkg.compress_bound(n, codec.prefs)
    d_comp = torch.empty(cap, dtype=torch.uint8, device="cuda")
    d_back = torch.empty(n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        c = codec.compress_device(d_in.data_ptr(), n, d_comp.data_ptr(), cap)
        sc = codec.stats()
        r, _ = codec.decompress_device(d_comp.data_ptr(), c, d_back.data_ptr(), n)
        sd = codec.stats()
        assert r == n
        return c, sc, sd

    for _ in range(args.warmup):
        csize, _, _ = step_device()
    assert torch.equal(d_back, d_in), "round trip is not bit-exact"

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t0 = time.perf_counter()
    ev = {"c_total": 0.0, "c_main": 0.0, "d_total": 0.0, "d_main": 0.0, "launches": 0}
    for _ in range(args.steps):
        csize, sc, sd = step_device()
        ev["c_total"] += sc.ms_total; ev["c_main"] += sc.ms_main
        ev["d_total"] += sd.ms_total; ev["d_main"] += sd.ms_main
        ev["launches"] += sc.kernel_launches + sd.kernel_launches
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    K = args.steps
    dev_ms = (ev["c_total"] + ev["d_total"]) / K             # CUDA events on the codec's own stream
    wall_ms = 1e3 * wall / K
    wall_ms, dev_ms, c_ms, d_ms = max_over_ranks([wall_ms, dev_ms, ev["c_total"] / K, ev["d_total"] / K], "cuda")

    # ---- end to end through the host-buffer C ABI: pinned host in/out, copies inside the timed region
    e2e = None
    if not args.skip_e2e:
        host_comp = torch.empty(cap, dtype=torch.uint8).pin_memory()
        host_back = torch.empty(n, dtype=torch.uint8).pin_memory()

        def step_host():
            c = codec.compress_host_ptr(host_in.data_ptr(), n, host_comp.data_ptr(), cap)
            r, _ = codec.decompress_host_ptr(host_comp.data_ptr(), c, host_back.data_ptr(), n)
            assert r == n
            return c
        for _ in range(2):
            step_host()
        assert torch.equal(host_back, host_in)
        barrier()
        t1 = time.perf_counter()
        ke = max(2, min(K, 5))
        for _ in range(ke):
            ce = step_host()
        barrier()
        e_ms = 1e3 * (time.perf_counter() - t1) / ke
        e_ms = max_over_ranks([e_ms], "cuda")[0]
        link = link_bandwidth(torch, min(n, 1 << 30), barrier)
        # host-link roofline of the round trip: compress moves U up and C down, decompress C up and U down, the two calls
        # run one after the other and each direction of the link is used by one transfer at a time
        bound_ms = 1e3 * (max(n / link["h2d"], ce / link["d2h"]) + max(ce / link["h2d"], n / link["d2h"])) / 1e9
        e2e = {"value": world * n / (e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": n + ce,
               "d2h_bytes_per_step": ce + n, "ms_per_step": e_ms, "steps": ke,
               "api": "b2lz4_compress_host + b2lz4_decompress_host (pinned host buffers)",
               "link": {"h2d_GBps": link["h2d"], "d2h_GBps": link["d2h"], "duplex_GBps_per_direction": link["duplex"],
                        "note": "pinned copies, all ranks at once, slowest rank", "bound_GBps": world * n / (bound_ms * 1e-3) / 1e9,
                        "frac": (world * n / (e_ms * 1e-3)) / (world * n / (bound_ms * 1e-3))}}

    # ---- a frame this codec did not write (rank 0 only; the other ranks wait at the barrier below)
    foreign = None
    if rank == 0 and not args.skip_foreign and args.workload == "loglike":
        try:
            foreign = foreign_leg(torch, pkg, codec, ref_lib(), host_in, d_in, n)
        except Exception as ex:                                   # reported, never fatal for the headline line
            foreign = {"error": repr(ex)}

    # ---- configs 3 and 5: ONE frame's blocks sharded over all GPUs of the job inside ONE process (rank 0); the other
    # ranks have freed their buffers and wait
    sharded = None
    if not args.skip_sharded and args.workload == "loglike":
        del d_comp, d_back
        if rank != 0:
            del d_in
        torch.cuda.empty_cache()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=side)
        if rank == 0:
            try:
                sharded = sharded_leg(torch, pkg, gen, world)
            except Exception as ex:
                sharded = {"error": repr(ex)}
        if world > 1:
            dist.barrier(group=side)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peaks()
    alg = n + csize                                           # U + C per direction
    rf_d = {"kernel": "decode_chains_kernel", "bound": "hbm", "achieved": alg / (ev["d_main"] / K * 1e-3) / 1e9,
            "peak": peak, "unit": "GB/s", "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg, "avg_launch_ms": ev["d_main"] / K}
    rf_c = {"kernel": "find_emit_kernel", "bound": "hbm", "achieved": alg / (ev["c_main"] / K * 1e-3) / 1e9,
            "peak": peak, "unit": "GB/s", "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg, "avg_launch_ms": ev["c_main"] / K}
    traffic = {}
    try:                                                      # measured DRAM bytes per launch (one ncu --set full capture per kernel)
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        c = tj.get("config", {})
        if (c.get("workload") == args.workload and c.get("size_mib") == args.size_mib and c.get("block_id") == args.block_id
                and c.get("sub_log2") == args.sub_log2):
            traffic = tj
    except Exception:
        pass
    for r in (rf_d, rf_c):
        r["frac"] = r["achieved"] / peak
        r["traffic"] = traffic.get(r["kernel"])
    dominant, other = (rf_c, rf_d) if ev["c_main"] >= ev["d_main"] else (rf_d, rf_c)

    cpu = None
    if not args.skip_cpu:
        lib = ref_lib()
        if lib is not None:
            sample = min(n, 1 << 30)                     # the whole stream of config 2: ~2.5 s of one core
            tc, td, cs = ref_roundtrip(lib, host_in.data_ptr(), sample, 1)
            cpu = {"value": sample / (tc + td) / 1e9, "unit": UNIT, "cores": 1, "kind": "reference",
                   "sample": f"first {sample >> 20} MiB of the same stream, reference conduit replay (Conduit.hsc call sequence on the unmodified lz4 library), 1 thread",
                   "compress_GBps": sample / tc / 1e9, "decompress_GBps": sample / td / 1e9, "ratio": sample / cs}

    line = {"metric": METRIC, "value": world * n / (wall_ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": wall_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload, "bytes_per_gpu": n, "block_size": block_size, "sub_block": 1 << min(args.sub_log2, 8 + 2 * args.block_id),
                       "frame": "seekable (B2IX index frame appended)" if index else "plain (drop-in default)",
                       "l2": "inputs larger than L2 (1 GiB stream vs 126 MB L2)", "parallelism": f"{world} independent streams, no collective"},
            "compress_GBps": world * n / (c_ms * 1e-3) / 1e9, "decompress_GBps": world * n / (d_ms * 1e-3) / 1e9,
            "device_ms_per_step": dev_ms, "ratio": n / csize, "compressed_bytes": csize,
            "roofline": dominant, "roofline_other": other, "cpu_baseline": cpu, "e2e": e2e, "foreign": foreign, "sharded": sharded,
            "gpu_launches": ev["launches"], "clocks": clocks}
    if cpu:
        line["ratio_vs_reference_default"] = (n / csize) / cpu["ratio"]
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
