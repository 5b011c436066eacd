"""what a concurrent D2H copy does to (a) a chain of small kernels, (b) the same chain as a CUDA graph, (c) one long kernel"""
import torch, json, time
n = 256 * 320 * 240
d = torch.zeros(n, dtype=torch.int16, device="cuda")
h = torch.empty(n, dtype=torch.int16).pin_memory()
cs = torch.cuda.Stream(); ws = torch.cuda.Stream()
x = torch.zeros(1 << 20, device="cuda"); big = torch.zeros(1 << 28, device="cuda")
def chain():
    for _ in range(60):
        x.add_(1.0)
def long():
    for _ in range(4):
        big.add_(1.0)
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(ws):
    chain(); torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=ws):
        chain()
def timed(fn, copies):
    torch.cuda.synchronize()
    if copies:
        with torch.cuda.stream(cs):
            for _ in range(30):
                h.copy_(d, non_blocking=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ws):
        fn()  # warm
        e0.record(ws)
        for _ in range(5):
            fn()
        e1.record(ws)
    e1.synchronize()
    t = e0.elapsed_time(e1) / 5
    torch.cuda.synchronize()
    return t
out = {}
for name, fn in (("chain60", chain), ("graph60", g.replay), ("long", long)):
    out[name] = {"alone_ms": timed(fn, False), "with_d2h_ms": timed(fn, True)}
print(json.dumps(out))
