"""Link ceiling for the e2e leg: concurrent pinned H2D (lists of one 4K frame, ~9.8 MB) and D2H
(one 4K Main10 picture, 24.9 MB) copies on two streams, no kernels.  Prints frames/s the bus
alone would allow, to compare with bench.py's e2e figure."""
import sys
import time
import torch

h2d_bytes = int(float(sys.argv[1])) if len(sys.argv) > 1 else 9_840_000
d2h_bytes = int(float(sys.argv[2])) if len(sys.argv) > 2 else 24_883_200
n = 200
dev = torch.device("cuda:0")
hin = [torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory() for _ in range(4)]
hout = [torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory() for _ in range(4)]
din = [torch.empty(h2d_bytes, dtype=torch.uint8, device=dev) for _ in range(4)]
dout = [torch.empty(d2h_bytes, dtype=torch.uint8, device=dev) for _ in range(4)]
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down):
    torch.cuda.synchronize()
    t = time.perf_counter()
    for i in range(n):
        if up:
            with torch.cuda.stream(s_in):
                din[i % 4].copy_(hin[i % 4], non_blocking=True)
        if down:
            with torch.cuda.stream(s_out):
                hout[i % 4].copy_(dout[i % 4], non_blocking=True)
    torch.cuda.synchronize()
    return time.perf_counter() - t


for name, up, down in (("h2d only", 1, 0), ("d2h only", 0, 1), ("both", 1, 1)):
    run(up, down)
    dt = run(up, down)
    gb = (up * h2d_bytes + down * d2h_bytes) * n / dt / 1e9
    print(f"{name}: {n / dt:8.1f} frames/s  {gb:6.1f} GB/s  (h2d {up * h2d_bytes * n / dt / 1e9:5.1f}, d2h {down * d2h_bytes * n / dt / 1e9:5.1f})")
