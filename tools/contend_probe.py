"""do device->host copies slow the stages down?  resident re-runs of the C2 batch with and without a stream of 39 MB D2H copies beside them"""
import importlib, sys, time, json, os
os.environ.setdefault('CUDA_DEVICE_MAX_CONNECTIONS', '32')
import numpy as np, torch
sys.path.insert(0, '.')
pkg = importlib.import_module("xash3d-fwgs_b200"); scene = importlib.import_module("xash3d-fwgs_b200.scene")
m = scene.make_map("small", ncameras=256)
r = scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
r.render_resident(m.cameras[:256])
n = 256 * 320 * 240
d = torch.zeros(n, dtype=torch.int16, device="cuda")
h = torch.empty(n, dtype=torch.int16).pin_memory()
hp = torch.ones(n, dtype=torch.int16).pin_memory()
s = torch.cuda.Stream()
out = {}
r.bench_resident(5)
out["alone"] = r.bench_resident(20)
for name, fn in (("with_d2h", lambda: h.copy_(d, non_blocking=True)), ("with_h2d", lambda: d.copy_(hp, non_blocking=True))):
    with torch.cuda.stream(s):
        for _ in range(40):
            fn()
    out[name] = r.bench_resident(20)
    s.synchronize()
print(json.dumps(out))
