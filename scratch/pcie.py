import torch, time
dev = torch.device("cuda", 0)
H, W = 1080, 1920
hm = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(16)]
ha = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(16)]
dm = torch.empty((16, H, W, 4), dtype=torch.uint8, device=dev)
da = torch.empty((16, 668, 1088, 3), dtype=torch.uint8, device=dev)
dd = torch.empty((16, H, W), dtype=torch.uint8, device=dev)
hd = [torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(16)]
def t(fn, n=10):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
def up_main():
    for i in range(16): dm[i].copy_(hm[i], non_blocking=True)
def up_add():
    for i in range(16): da[i].copy_(ha[i][:668, :1088], non_blocking=True)
def down():
    for i in range(16): hd[i].copy_(dd[i], non_blocking=True)
a = t(up_main); b = t(up_add); c = t(down)
print(f"H2D main 133 MB: {a:.3f} ms = {132.7/a:.1f} GB/s;  H2D add (2D) 35 MB: {b:.3f} ms = {34.9/b:.1f} GB/s;  D2H 33 MB: {c:.3f} ms = {33.2/c:.1f} GB/s")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def both():
    with torch.cuda.stream(s1):
        up_main(); up_add()
    with torch.cuda.stream(s2):
        down()
print("H2D + D2H concurrently:", round(t(both), 3), "ms")
