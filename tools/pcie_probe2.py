"""Do uploads in frame-sized pieces overlap with downloads?  (torch streams, pinned memory)"""
import torch, time
sizes_up = [810520, 16320, 420, 2691720, 17418240, 16320, 519901, 519901, 129600, 4080, 81600, 4080]
sizes_dn = [8192 * 2160, 4352 * 1080, 4352 * 1080]
NS = 8
hu = [[torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes_up] for _ in range(NS)]
du = [[torch.empty(n, dtype=torch.uint8, device="cuda") for n in sizes_up] for _ in range(NS)]
hd = [[torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes_dn] for _ in range(NS)]
dd = [[torch.empty(n, dtype=torch.uint8, device="cuda") for n in sizes_dn] for _ in range(NS)]
su = [torch.cuda.Stream() for _ in range(NS)]
sd = torch.cuda.Stream()
def run(up, dn, frames=48):
    torch.cuda.synchronize(); t = time.perf_counter()
    for f in range(frames):
        k = f % NS
        if up:
            with torch.cuda.stream(su[k]):
                for a, b in zip(du[k], hu[k]): a.copy_(b, non_blocking=True)
        if dn:
            with torch.cuda.stream(sd):
                for a, b in zip(hd[k], dd[k]): a.copy_(b, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / frames * 1e3
run(True, True, 8)
print("upload pieces only   %.3f ms/frame" % run(True, False))
print("download only        %.3f ms/frame" % run(False, True))
print("both                 %.3f ms/frame" % run(True, True))
