"""Generates the committed golden fixtures.  Run HERE (needs /root/reference and the oracle build):

    python tests/golden/make_golden.py

1. gamma_golden.npz - the known-answer arrays of the reference's own Test_PrecomputedGammaTables
   (engine/client/gamma.c:245-268, data at :281.. and :357..), parsed from the source text.  They pin the
   host's restatement of BuildGammaTable, whose lightgammatable feeds R_BuildLightMap (ref/soft/r_surf.c:216).
2. frames_golden.npz / frames_golden.json - outputs of the UNMODIFIED ref_soft (oracle/_ref/libref_soft.so)
   driven by the headless host on the seeded synthetic maps: full frames for the tiny map, SHA-256 of every
   frame and of every surface-cache block for the tiny and small maps.  Derived data, labelled as such: the
   reference ships no image/span/cache fixtures.
"""
import hashlib
import importlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("REF", "/root/reference")


def parse_gamma():
    src = open(os.path.join(REF, "engine/client/gamma.c")).read()
    src = src[src.index("precomputed_gamma_tables_t *Test_GetGammaTables( int i )\n{"):]
    sets = []
    for blk in re.split(r"(?=\.gamma = )", src)[1:]:
        if ".gamma" not in blk:
            continue
        d = {}
        for k in ("gamma", "brightness", "texgamma", "lightgamma"):
            d[k] = float(re.search(r"\.%s = ([0-9.]+)" % k, blk).group(1))
        for k in ("texgammatable", "lightgammatable", "lineargammatable", "screengammatable"):
            body = re.search(r"\.%s = \{(.*?)\}" % k, blk, re.S).group(1)
            d[k] = np.array([int(x) for x in re.findall(r"-?\d+", body)], np.int32)
        assert len(d["texgammatable"]) == 256 and all(len(d[k]) == 1024 for k in ("lightgammatable", "lineargammatable", "screengammatable"))
        sets.append(d)
    assert len(sets) == 2, len(sets)
    out = {}
    for i, d in enumerate(sets):
        out[f"params{i}"] = np.array([d["gamma"], d["brightness"], d["texgamma"], d["lightgamma"]], np.float32)
        for k in ("texgammatable", "lightgammatable", "lineargammatable", "screengammatable"):
            out[f"{k}{i}"] = d[k]
    np.savez_compressed(os.path.join(HERE, "gamma_golden.npz"), **out)
    print("gamma_golden.npz:", [(k, v.shape) for k, v in out.items()][:5])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def frames():
    pkg = importlib.import_module("xash3d-fwgs_b200")
    scene = importlib.import_module("xash3d-fwgs_b200.scene")
    meta = {"note": "derived from oracle/_ref/libref_soft.so (unmodified ref_soft, gcc -O2) on seeded synthetic maps", "cases": {}}
    store = {}
    for which, n, (w, h) in (("tiny", 8, (320, 240)), ("small", 16, (320, 240)), ("small", 2, (640, 480))):
        m = scene.make_map(which, ncameras=max(n, 16))
        r = scene.Renderer(pkg.REF_SOFT_PATH, w, h, m)
        col, z = r.render_views(m.cameras[:n])
        blocks = {}
        nsurf = r.host.world_info()["numsurfaces"]
        for s in range(nsurf):
            for mip in range(4):
                b = r.cache_block(s, mip)
                if b is not None:
                    blocks[f"{s}.{mip}"] = sha(b)
        key = f"{which}_{n}x{w}x{h}"
        meta["cases"][key] = {
            "map": which, "views": n, "width": w, "height": h, "bsp_sha256": hashlib.sha256(m.bsp).hexdigest(),
            "colour_sha256": [sha(col[i]) for i in range(n)], "z_sha256": [sha(z[i]) for i in range(n)],
            "cache_blocks_sha256": blocks,
        }
        if which == "tiny":
            store["tiny_colour"] = col
            store["tiny_z"] = z
        print(key, "frames", n, "cache blocks", len(blocks))
    np.savez_compressed(os.path.join(HERE, "frames_golden.npz"), **store)
    json.dump(meta, open(os.path.join(HERE, "frames_golden.json"), "w"), indent=0, sort_keys=True)


if __name__ == "__main__":
    parse_gamma()
    frames()
