"""H2D bandwidth of this box for the e2e path: pinned vs write-combined pinned, one vs several copy streams, NUMA."""
import os, time, sys
import torch
from cuda import cudart

def chk(r):
    if r[0] != cudart.cudaError_t.cudaSuccess: raise RuntimeError(str(r[0]))
    return r[1] if len(r) == 2 else r[1:]

N = 185 * 1024 * 1024
torch.cuda.init(); torch.zeros(1, device="cuda")
try:
    import pynvml
    pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
    pynvml.nvmlDeviceSetCpuAffinity(h); print("affinity", len(os.sched_getaffinity(0)), "cpus")
except Exception as e:
    print("no affinity", e)
d = chk(cudart.cudaMalloc(N))
streams = [chk(cudart.cudaStreamCreate()) for _ in range(4)]
for name, flags in (("default", cudart.cudaHostAllocDefault), ("writecombined", cudart.cudaHostAllocWriteCombined)):
    hp = chk(cudart.cudaHostAlloc(N, flags))
    t = time.time()
    import ctypes; ctypes.memset(hp, 7, N)
    fill = time.time() - t
    for ns in (1, 2, 4):
        part = N // ns
        for rep in range(2):
            cudart.cudaDeviceSynchronize(); t = time.time()
            for it in range(10):
                for s in range(ns):
                    cudart.cudaMemcpyAsync(d + s * part, hp + s * part, part, cudart.cudaMemcpyKind.cudaMemcpyHostToDevice, streams[s])
            cudart.cudaDeviceSynchronize(); dt = (time.time() - t) / 10
        print(f"{name:14s} streams={ns} {dt*1e3:.2f} ms {N/dt/1e9:.1f} GB/s (host fill {N/fill/1e9:.1f} GB/s)")
    cudart.cudaFreeHost(hp)
