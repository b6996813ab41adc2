"""PCIe copy-rate probe (pinned host memory): the ceiling for the e2e number."""
import torch, time
n = 240 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(64 << 20, dtype=torch.uint8).pin_memory(); d2 = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
def t(f, reps=5):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
a = t(lambda: d.copy_(h, non_blocking=True)); print("H2D 240MB: %.2f ms  %.1f GB/s" % (a * 1e3, n / a / 1e9))
b = t(lambda: h2.copy_(d2, non_blocking=True)); print("D2H 64MB: %.2f ms  %.1f GB/s" % (b * 1e3, (64 << 20) / b / 1e9))
s1 = torch.cuda.Stream(); s2 = torch.cuda.Stream()
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
c = t(both); print("both concurrently: %.2f ms" % (c * 1e3))
# chunked uploads interleaved with a small kernel on the same stream (the shape of orb_stereo_frames' upload stream)
import numpy as np
hn = np.random.randint(0, 255, (512, 376, 1241), dtype=np.uint8)
hp = torch.from_numpy(hn).pin_memory()
dd = torch.empty((512, 376, 1241), dtype=torch.uint8, device="cuda"); d3 = torch.empty((32, 376, 1280), dtype=torch.uint8, device="cuda")
def chunks(k=32, kern=True):
    with torch.cuda.stream(s1):
        for c in range(0, 512, k):
            dd[c:c + k].copy_(hp[c:c + k], non_blocking=True)
            if kern: d3[:, :, :1241].copy_(dd[c:c + 32])
for kern in (False, True):
    e = t(lambda: chunks(32, kern)); print("16 x 15MB chunks, kernel between=%s: %.2f ms  %.1f GB/s" % (kern, e * 1e3, hp.numel() / e / 1e9))
