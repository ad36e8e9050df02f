import torch,time
h=torch.empty(185*1024*1024,dtype=torch.uint8,pin_memory=True); d=torch.empty_like(h,device='cuda')
h2=torch.empty(38*1024*1024,dtype=torch.uint8,pin_memory=True); d2=torch.empty_like(h2,device='cuda')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream()
for name,fn in (("h2d",lambda: d.copy_(h,non_blocking=True)),("d2h",lambda: h2.copy_(d2,non_blocking=True))):
    torch.cuda.synchronize(); t=time.time()
    for _ in range(10): fn()
    torch.cuda.synchronize(); dt=(time.time()-t)/10
    print(name, "%.2f ms"%(dt*1e3), "%.1f GB/s"%((h.numel() if name=="h2d" else h2.numel())/dt/1e9))
torch.cuda.synchronize(); t=time.time()
for _ in range(10):
    with torch.cuda.stream(s1): d.copy_(h,non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
torch.cuda.synchronize(); print("both %.2f ms"%((time.time()-t)/10*1e3))
import subprocess; print(subprocess.run("nvidia-smi -q | grep -A6 'GPU Link Info' | head -12",shell=True,capture_output=True,text=True).stdout)
