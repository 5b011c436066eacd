"""ad-hoc GPU check: parity of libref_cuda.so vs the oracle + stage timings (development aid)."""
import sys, importlib, time, json
sys.path.insert(0, '.')
import numpy as np
pkg = importlib.import_module('xash3d-fwgs_b200')
scene = importlib.import_module('xash3d-fwgs_b200.scene')
which = sys.argv[1] if len(sys.argv) > 1 else 'small'
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
W = int(sys.argv[3]) if len(sys.argv) > 3 else 320
H = int(sys.argv[4]) if len(sys.argv) > 4 else 240
m = scene.make_map(which, ncameras=n)
print('map', m.stats)
cu = scene.Renderer(pkg.REF_CUDA_PATH, W, H, m)
print('device', cu.device_name())
t = time.time(); cc, cz = cu.render_views(m.cameras[:n]); print('cuda first batch s', time.time() - t)
print(cu.counters())
so = scene.Renderer(pkg.REF_SOFT_PATH, W, H, m, cvars={'sw_maxsurfs': 60000, 'sw_maxedges': 250000})  # ref_soft drops faces past 2000 / edges past 4000 by default
rc, rz = so.render_views(m.cameras[:n]); print('soft s', so.last_seconds)
print('colour same', (rc == cc).mean(), 'z same', (rz == cz).mean())
bad = [(i, float((rc[i] == cc[i]).mean())) for i in range(n) if (rc[i] != cc[i]).any()]
print('views with diffs', bad[:10])
for it in range(2):
    tot, st = cu.bench_resident(10)
    print('resident ms/batch', tot, 'stages', st, 'Mpix/s', n * W * H / tot / 1e3)
tot, st = cu.bench_resident(10, True)
print('rebuild-cache ms/batch', tot, 'stages', st)
print('launches', cu.launch_count())
