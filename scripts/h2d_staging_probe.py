"""Probe: host-to-device copy of one [8192, 1024] f32 latent batch from pageable, pinned and staged-through-pinned memory (what
bounds the e2e leg of bench.py on the host side)."""
import torch, time, numpy as np
torch.cuda.init()
x = torch.from_numpy(np.random.randn(8192,1024).astype(np.float32))
pin = torch.empty_like(x).pin_memory()
dev = torch.empty_like(x, device="cuda")
print("threads", torch.get_num_threads())
def t(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/n*1e3
print("pageable -> device  %.3f ms" % t(lambda: dev.copy_(x, non_blocking=True)))
print("pinned -> device    %.3f ms" % t(lambda: dev.copy_(pin, non_blocking=True)))
print("pageable -> pinned  %.3f ms" % t(lambda: pin.copy_(x)))
def staged():
    pin.copy_(x); dev.copy_(pin, non_blocking=True)
print("staged whole        %.3f ms" % t(staged))
def staged4():
    for c in range(4):
        s=slice(c*2048,(c+1)*2048)
        pin[s].copy_(x[s]); dev[s].copy_(pin[s], non_blocking=True)
print("staged 4 chunks     %.3f ms" % t(staged4))
for nt in (4,8,16):
    torch.set_num_threads(nt)
    print(nt, "threads: pageable -> pinned %.3f ms, staged4 %.3f ms" % (t(lambda: pin.copy_(x)), t(staged4)))
