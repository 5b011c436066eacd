import importlib, sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, '..')
pkg = importlib.import_module("xash3d-fwgs_b200"); scene = importlib.import_module("xash3d-fwgs_b200.scene")
which = sys.argv[1] if len(sys.argv) > 1 else "tiny"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4
w, h = (1920, 1080) if which == "large" else (320, 240)
m = scene.make_map(which, ncameras=max(n, 16))
cu = scene.Renderer(pkg.REF_CUDA_PATH, w, h, m)
a, az = cu.render_views(m.cameras[:n])
print("rendered", which, n, cu.counters())
so = scene.Renderer(pkg.REF_SOFT_PATH, w, h, m, cvars={"sw_maxsurfs": 40000, "sw_maxedges": 160000})
rc, rz = so.render_views(m.cameras[:n])
print("identical colour", (a == rc).mean(), "z", (az == rz).mean())
