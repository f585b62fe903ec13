"""GPU box, under torchrun: pinned-copy bandwidth of every rank's host link with all ranks copying at once, first with the
inherited CPU affinity, then with the process bound to the CPUs NVML reports as local to its GPU (pinned buffers are
allocated after binding, so first touch places them on that node)."""
import os, sys, time
import torch, torch.distributed as dist
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", local))
def barrier():
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
def local_cpus():
    import pynvml
    pynvml.nvmlInit()
    p = torch.cuda.get_device_properties(local)
    bus = f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
    n = os.cpu_count(); words = (n + 63) // 64
    m = pynvml.nvmlDeviceGetCpuAffinity(h, words)
    return [i for i in range(n) if (m[i // 64] >> (i % 64)) & 1]
def measure(tag):
    n = 1 << 30
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); h.zero_()
    h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2.zero_()
    d = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    s2 = torch.cuda.Stream()
    res = {}
    for name in ("h2d", "d2h", "both"):
        barrier(); t0 = time.perf_counter()
        for _ in range(4):
            if name in ("h2d", "both"): d.copy_(h, non_blocking=True)
            if name in ("d2h", "both"):
                with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        res[name] = 4 * n / dt / 1e9
    print(f"rank {rank} {tag}: h2d {res['h2d']:.1f} d2h {res['d2h']:.1f} both(each dir) {res['both']:.1f} GB/s  cpus {sorted(os.sched_getaffinity(0))[:4]}..({len(os.sched_getaffinity(0))})", flush=True)
print(f"rank {rank}: cpu_count {os.cpu_count()} inherited affinity {len(os.sched_getaffinity(0))} cpus", flush=True)
measure("inherited")
try:
    c = local_cpus()
    print(f"rank {rank}: NVML local cpus {c[:3]}..{c[-3:]} ({len(c)})", flush=True)
    allowed = sorted(set(c) & os.sched_getaffinity(0))
    if allowed: os.sched_setaffinity(0, allowed)
    measure("bound")
except Exception as e:
    print(f"rank {rank}: binding failed: {e!r}", flush=True)
if world > 1: dist.destroy_process_group()
