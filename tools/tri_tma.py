import torch, triton, triton.language as tl, os
def alloc_fn(size, align, stream): return torch.empty(size, device="cuda", dtype=torch.int8)
triton.set_allocator(alloc_fn)
@triton.jit
def k(a_ptr, o_ptr, M, N, BM: tl.constexpr, BN: tl.constexpr):
    d = tl.make_tensor_descriptor(a_ptr, shape=[M, N], strides=[N, 1], block_shape=[BM, BN])
    x = d.load([7, 6])
    offs = tl.arange(0, BM)[:, None] * BN + tl.arange(0, BN)[None, :]
    tl.store(o_ptr + offs, x)
a = torch.arange(2320*4096, device="cuda", dtype=torch.int32).reshape(2320, 4096).to(torch.int16)
o = torch.zeros(16*16, device="cuda", dtype=torch.int16)
h = k[(1,)](a, o, 2320, 4096, 16, 16)
torch.cuda.synchronize()
print("triton tma ok:", torch.equal(o.reshape(16,16), a[7:23, 6:22]))
open("gpurun_out/tri_tma.ptx","w").write(h.asm["ptx"])
try:
    open("gpurun_out/tri_tma.sass","w").write(h.asm["sass"])
except Exception as e: print("no sass", e)
