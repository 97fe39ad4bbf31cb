"""The e2e leg's link ceiling with N GPUs at once: N processes (one per GPU) run tools/pcie_ceiling.py's "both" pattern
(pinned 9.8 MB up + 24.9 MB down per frame, two streams, no kernels) at the same time; prints per-GPU and aggregate frame/s.
usage: python tools/pcie_ceiling_multi.py N"""
import subprocess
import sys
import time

CHILD = r'''
import sys, time, torch
dev = torch.device(f"cuda:{sys.argv[1]}"); torch.cuda.set_device(dev)
h2d, d2h, n = 9_840_000, 24_883_200, 300
hin = [torch.empty(h2d, dtype=torch.uint8).pin_memory() for _ in range(4)]
hout = [torch.empty(d2h, dtype=torch.uint8).pin_memory() for _ in range(4)]
din = [torch.empty(h2d, dtype=torch.uint8, device=dev) for _ in range(4)]
dout = [torch.empty(d2h, dtype=torch.uint8, device=dev) for _ in range(4)]
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
def run():
    torch.cuda.synchronize(); t = time.perf_counter()
    for i in range(n):
        with torch.cuda.stream(s_in): din[i % 4].copy_(hin[i % 4], non_blocking=True)
        with torch.cuda.stream(s_out): hout[i % 4].copy_(dout[i % 4], non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t
run()
while time.time() < float(sys.argv[2]): time.sleep(0.001)      # common start
dt = run()
print(f"{n / dt:.1f}")
'''
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
start = time.time() + 25
procs = [subprocess.Popen([sys.executable, "-c", CHILD, str(i), str(start)], stdout=subprocess.PIPE, text=True) for i in range(n)]
fps = [float(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
print(f"{n} GPUs at once: per GPU {[round(f) for f in fps]} frame/s, aggregate {sum(fps):.0f} frame/s = {sum(fps) * 34.72e6 / 1e9:.1f} GB/s combined")
