"""device->host copy rate of the box for a C2 batch's frames (39.3 MB) into pinned memory: the floor of the e2e leg"""
import torch, time, json
n = 256 * 320 * 240
d = torch.zeros(n, dtype=torch.int16, device="cuda")
h = torch.empty(n, dtype=torch.int16).pin_memory()
s = torch.cuda.Stream()
out = {}
for chunks in (1, 4):
    with torch.cuda.stream(s):
        for _ in range(5):
            h.copy_(d, non_blocking=True)
        s.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            for c in range(chunks):
                a, b = n * c // chunks, n * (c + 1) // chunks
                h[a:b].copy_(d[a:b], non_blocking=True)
        s.synchronize()
        dt = (time.perf_counter() - t0) / 50
    out[f"chunks{chunks}"] = {"ms": dt * 1e3, "GBps": n * 2 / dt / 1e9}
print(json.dumps(out))
