"""Where do the pinned buffers of the ranks live, and what does the aggregate H2D bandwidth look like? (torchrun, N ranks)"""
import os, re, sys, time
import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
import pynvml
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(lr)
before = len(os.sched_getaffinity(0))
pynvml.nvmlDeviceSetCpuAffinity(h)
aff = sorted(os.sched_getaffinity(0))
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))

def node_of(t):
    addr = t.data_ptr()
    out = {}
    for line in open("/proc/self/numa_maps"):
        a = int(line.split()[0], 16)
        if a <= addr < a + (1 << 34) and re.search(r"N\d+=", line):
            cand = (addr - a, line)
            if "best" not in out or cand[0] < out["best"][0]:
                out["best"] = cand
    if "best" not in out: return "?"
    return " ".join(re.findall(r"N\d+=\d+", out["best"][1]))[:60]

N = 185 * 1024 * 1024
hbuf = torch.empty(N, dtype=torch.uint8, pin_memory=True); hbuf.fill_(3)
dbuf = torch.empty(N, dtype=torch.uint8, device="cuda")
pci = pynvml.nvmlDeviceGetPciInfo(h).busId
msg = f"rank {rank} gpu {lr} pci {pci} cpus {before}->{len(aff)} [{aff[0]}..{aff[-1]}] pinned pages: {node_of(hbuf)}"
gathered = [None] * world
dist.all_gather_object(gathered, msg)
if rank == 0:
    print("\n".join(gathered)); 
    import subprocess
    print(subprocess.run("nvidia-smi topo -m | head -14; lscpu | grep -i -E 'numa|model name|socket'", shell=True, capture_output=True, text=True).stdout)

def bw(active):
    dist.barrier(); torch.cuda.synchronize(); t = time.perf_counter()
    if active:
        for _ in range(10): dbuf.copy_(hbuf, non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    dist.barrier()
    return N * 10 / dt / 1e9 if active else 0.0

for label, ranks in (("all", range(world)), ("rank0 only", [0]), ("first half", range(world // 2)), ("second half", range(world // 2, world)), ("even", range(0, world, 2))):
    v = bw(rank in ranks)
    t = torch.tensor([v], device="cuda"); lst = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(lst, t)
    if rank == 0:
        vals = [round(float(x), 1) for x in lst]
        print(label, "GB/s per rank", vals, "sum", round(sum(vals), 1))
dist.destroy_process_group()
