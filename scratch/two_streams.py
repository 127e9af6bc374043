"""experiment: does running two half-batches on two streams (two contexts = two scratch sets) beat one stream?"""
import sys, torch
sys.path.insert(0, ".")
import stepth_b200 as sb
from stepth_b200 import device, synth
W, H, D, P = 1920, 1080, 128, [36, 36, 36]
dev = torch.device("cuda", 0)
kind = sys.argv[1] if len(sys.argv) > 1 else "nat"
ms, as_ = [], []
for i in range(16):
    m, a = synth.make_pair(kind, W, H, seed=i + 1, device=dev); ms.append(m); as_.append(a)
mains, adds = torch.stack(ms).contiguous(), torch.stack(as_).contiguous()
def bench(nstreams, steps=10):
    n = 16 // nstreams
    ctxs = [sb.Context(0) for _ in range(nstreams)]
    pipes = [device.DensePipeline(c, n, W, H, W, H, P, D, sb.RESAMPLE_U8) for c in ctxs]
    streams = [torch.cuda.Stream() for _ in range(nstreams)]
    def step():
        for k in range(nstreams):
            with torch.cuda.stream(streams[k]):
                pipes[k].run(mains[k * n:(k + 1) * n], adds[k * n:(k + 1) * n])
    for _ in range(3): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams: s.wait_event(e0)
    for _ in range(steps): step()
    for s in streams: torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    ref = torch.cat([p.depth for p in pipes])
    return e0.elapsed_time(e1) / steps, ref
t1, r1 = bench(1)
for ns in (2, 4):
    t, r = bench(ns)
    print(kind, "streams", ns, "ms/step", round(t, 3), "vs 1 stream", round(t1, 3), "equal", bool(torch.equal(r, r1)))
