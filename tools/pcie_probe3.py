"""Does the number of copies per frame matter?  Uploads of a sparse-coefficient 4K frame as the
13 lists the library pushes, or as one piece of the same size; downloads as 3 planes or 1 piece.
(torch streams, pinned memory; uploads on 8 streams in turn, downloads on one stream)"""
import torch, time
lists = [960000, 16320, 420, 2691720, 5037296, 16320, 16320, 519901, 519901, 129600, 4080, 81600, 4080]
planes = [7680 * 2160, 3840 * 1080, 3840 * 1080]
NS = 8
def bufs(sizes):
    return ([[torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes] for _ in range(NS)],
            [[torch.empty(n, dtype=torch.uint8, device="cuda") for n in sizes] for _ in range(NS)])
su = [torch.cuda.Stream() for _ in range(NS)]
sd = torch.cuda.Stream()
def run(up, dn, frames=64):
    torch.cuda.synchronize(); t = time.perf_counter()
    for f in range(frames):
        k = f % NS
        if up:
            with torch.cuda.stream(su[k]):
                for a, b in zip(up[1][k], up[0][k]): a.copy_(b, non_blocking=True)
        if dn:
            with torch.cuda.stream(sd):
                for a, b in zip(dn[0][k], dn[1][k]): a.copy_(b, non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / frames * 1e3
up13, up1, up4 = bufs(lists), bufs([sum(lists)]), bufs([lists[0], lists[3], lists[4], sum(lists) - lists[0] - lists[3] - lists[4]])
dn3, dn1 = bufs(planes), bufs([sum(planes)])
run(up13, dn3, 16)
for name, u, d in [("up 13 lists only", up13, None), ("up 1 piece only", up1, None), ("down 3 planes only", None, dn3),
                   ("down 1 piece only", None, dn1), ("13 lists + 3 planes", up13, dn3), ("4 pieces + 3 planes", up4, dn3),
                   ("1 piece + 3 planes", up1, dn3), ("1 piece + 1 piece", up1, dn1), ("13 lists + 1 piece", up13, dn1)]:
    ms = min(run(u, d) for _ in range(3))
    mb = ((sum(lists) if u else 0) + (sum(planes) if d else 0)) / 1e6
    print(f"{name:22s} {ms:.3f} ms/frame  {mb / ms:.1f} GB/s  -> {1e3 / ms:.0f} frame/s")
