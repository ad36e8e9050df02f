"""Distribution cost against corner density: the bench image plus Gaussian noise of growing sigma."""
import os, sys
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", "..", "snake-slam_b200")]
import numpy as np, torch
import snakeorb
from snakeorb import synthetic as S

B, w, h = 128, 752, 480
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
for sigma in (0, 3, 5, 8, 12):
    rng = np.random.default_rng(sigma)
    imgs = np.stack([np.clip(S.image(100 + i, w, h).astype(np.float32) + rng.normal(0, sigma, (h, w)), 0, 255).astype(np.uint8) for i in range(8)])
    batch = np.stack([imgs[i % 8] for i in range(B)])
    dev = torch.from_numpy(batch).cuda()
    ex = snakeorb.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=B)
    ex.set_stream(st.cuda_stream)
    for _ in range(3): ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
    ex.DetectBatch(batch[:2])
    per_level = [len(ex.candidates(l, 0)) for l in range(8)]
    ex.set_profiling(True); ex.stage_ms(reset=True)
    for _ in range(5): ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
    torch.cuda.synchronize()
    sm, _ = ex.stage_ms(reset=True)
    print("sigma", sigma, "candidates/img", sum(per_level), "level 0:", per_level[0], "fast / distribute ms per 128 images:", round(float(sm[2]) / 5, 3), round(float(sm[3]) / 5, 3))
