import sys, importlib
sys.path.insert(0, '.')
import numpy as np
pkg = importlib.import_module('xash3d-fwgs_b200'); scene = importlib.import_module('xash3d-fwgs_b200.scene')
m = scene.make_map('small', ncameras=256)
cu = scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
cu.render_views(m.cameras[:256])
cu.render_views(m.cameras[:256])
a = np.zeros(64, np.uint64); cu._counters(a.ctypes.data, 64)
nl = int(a[20])
print('lines', nl, 'sort mean cycles/line', int(a[16]) / nl, 'max', int(a[17]), 'replay mean', int(a[18]) / nl, 'max', int(a[19]), 'warp total max cycles', int(a[21]))
print('events total', int(a[24]), 'max/line', int(a[25]))
print(cu.counters())
for i, name in enumerate(['stage', 'fatpvs', 'markleaves', 'P1', 'P2', 'P3', 'P4', 'P5']):
    print('vis', name, 'mean cycles/block', int(a[32 + 2 * i]) / 256, 'max', int(a[33 + 2 * i]))
nb = 256 * 8
for i, name in enumerate(['gather', 'keys', 'sort', 'localnum', 'count', 'fill']):
    print('events', name, 'mean cycles/block', int(a[48 + i]) / nb, 'max', int(a[56 + i]))
