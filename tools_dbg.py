import sys, importlib
sys.path.insert(0, '.')
import numpy as np
pkg = importlib.import_module('xash3d-fwgs_b200'); scene = importlib.import_module('xash3d-fwgs_b200.scene')
m = scene.make_map('small', ncameras=256)
cu = scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
cu.render_views(m.cameras[:256])
cu.render_views(m.cameras[:256])
a = np.zeros(32, np.uint64); cu._counters(a.ctypes.data, 32)
for i, name in enumerate(['count', 'fill', 'sort', 'replay']):
    print(name, 'mean cycles/block', int(a[16 + 2 * i]) / 1024, 'max', int(a[17 + 2 * i]))
