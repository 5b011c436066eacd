"""profiling driver: one C2 batch (or --config) rendered, then re-run resident a few times (ncu attaches to this)"""
import importlib, sys
sys.path.insert(0, '.'); sys.path.insert(0, '..')
pkg = importlib.import_module("xash3d-fwgs_b200"); scene = importlib.import_module("xash3d-fwgs_b200.scene")
cfg = sys.argv[1] if len(sys.argv) > 1 else "c2"
n, w, h, which, kw = {"c2": (256, 320, 240, "small", {}), "c3": (8, 1920, 1080, "large", {}), "c5": (1024, 320, 240, "small", {"water_prob": (1, 5)}),
                       "c4": (256, 320, 240, "small", {})}[cfg]
m = scene.make_map(which, ncameras=n, **kw)
cu = scene.Renderer(pkg.REF_CUDA_PATH, w, h, m)
if cfg == "c4":
    cu.set_time(1.0); scene.add_studio_models(cu, m.cameras, 64)
if cfg == "c5":
    scene.add_sprites(cu, m.cameras, 32)
cu.render_resident(m.cameras[:n])
tot, st = cu.bench_resident(3, flush_l2=True, rebuild_cache=(cfg == "c3" or "rebuild" in sys.argv))
print(cfg, "ms", tot, st, cu.counters())
