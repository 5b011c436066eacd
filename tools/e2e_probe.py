"""where the pipelined e2e step goes: host time inside RefCuda_RenderViewsAsync vs the step period"""
import importlib, sys, time, json, os
os.environ.setdefault('CUDA_DEVICE_MAX_CONNECTIONS', '32')
import numpy as np, torch
sys.path.insert(0, '.')
pkg = importlib.import_module("xash3d-fwgs_b200"); scene = importlib.import_module("xash3d-fwgs_b200.scene")
m = scene.make_map("small", ncameras=256)
r = scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
n = 256
pins = [torch.empty((n, 240, 320), dtype=torch.int16).pin_memory() for _ in range(2)]
vp = r.viewpasses(m.cameras[:n])
out = {}
for mode in ("sync", "async"):
    for k in range(5):
        r.render_prepared(vp, pins[0].data_ptr(), None)
    calls = []
    t0 = time.perf_counter()
    for k in range(200):
        c0 = time.perf_counter()
        if mode == "sync":
            r.render_prepared(vp, pins[k & 1].data_ptr(), None)
        else:
            r.render_prepared_async(vp, pins[k & 1].data_ptr(), None)
        calls.append(time.perf_counter() - c0)
    r.wait()
    dt = (time.perf_counter() - t0) / 200
    out[mode] = {"step_ms": dt * 1e3, "call_ms_mean": float(np.mean(calls)) * 1e3, "call_ms_min": float(np.min(calls)) * 1e3}
# host-only part: resident batch (no outputs) call time
t0 = time.perf_counter()
for k in range(200):
    r.render_prepared(vp, None, None)
out["resident_call_ms"] = (time.perf_counter() - t0) / 200 * 1e3
print(json.dumps(out))
