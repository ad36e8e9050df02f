"""Throughput of the extractor on adversarial inputs (pure noise: every pixel a FAST candidate; constant: none)."""
import os, sys, time
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", "..", "snake-slam_b200")]
import numpy as np, torch
import snakeorb
from snakeorb import synthetic as S

B, w, h = 128, 752, 480
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
for kind in ("image", "noise", "constant", "checker"):
    if kind == "image":
        imgs = np.stack([S.image(100 + i, w, h) for i in range(8)])
    else:
        imgs = np.stack([S.adversarial(kind, w, h, 5 + i) if kind == "noise" else S.adversarial(kind, w, h) for i in range(8)])
    batch = np.stack([imgs[i % 8] for i in range(B)])
    dev = torch.from_numpy(batch).cuda()
    ex = snakeorb.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=B)
    ex.set_stream(st.cuda_stream)
    for _ in range(3): ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(10): ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    res = ex.DetectBatch(batch[:2])
    ncand = sum(len(ex.candidates(l, 0)) for l in range(8))
    ex.set_profiling(True); ex.stage_ms(reset=True)
    for _ in range(5): ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
    torch.cuda.synchronize()
    sm, _ = ex.stage_ms(reset=True)
    print(kind, "images/s", round(B / ms * 1e3), "kp/img", len(res[0][0]), "raw candidates/img", ncand, "stage ms per 128 images:", [round(float(x) / 5, 3) for x in sm[1:6]])
