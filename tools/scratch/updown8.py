"""What the host side of the box allows when every GPU uploads 92 MB chunks AND downloads 19 MB result chunks at the same
time (the traffic of bench.py's end-to-end path without any kernel): run under torchrun with one rank per GPU."""
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
dist.init_process_group("gloo")
torch.cuda.set_device(rank)
up_h = torch.empty(92 << 20, dtype=torch.uint8, pin_memory=True)
up_d = torch.empty(92 << 20, dtype=torch.uint8, device="cuda")
dn_d = torch.empty(19 << 20, dtype=torch.uint8, device="cuda")
dn_h = torch.empty(19 << 20, dtype=torch.uint8, pin_memory=True)
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("upload only", "upload + download"):
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(20):
        with torch.cuda.stream(s_up):
            up_d.copy_(up_h, non_blocking=True)
        if mode != "upload only":
            with torch.cuda.stream(s_dn):
                dn_h.copy_(dn_d, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = torch.tensor([20 * (92 << 20) / dt / 1e9], dtype=torch.float64)
    allr = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(allr, gbs)
    if rank == 0:
        r = [round(float(x), 1) for x in allr]
        print(mode, "GB/s per GPU", r, "sum", round(sum(r), 1), "=> pairs/s ceiling", round(sum(r) * 1e9 / (2 * 752 * 480)))
dist.destroy_process_group()
