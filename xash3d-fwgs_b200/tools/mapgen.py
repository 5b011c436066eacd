"""Procedural BSP30 (Half-Life) map generator: synthetic worlds for the ref_cuda benches and parity tests.

Half-Life content is unavailable offline, so every asset is synthetic (BASELINE.json north_star).
The output is a byte-exact BSP version 30 file image as laid out in common/bspfile.h:92-321 of the
reference: lumps ENTITIES, PLANES, TEXTURES (embedded 8-bit miptex + palette, common/wadfile.h:79),
VERTEXES, VISIBILITY (zero-run RLE rows), NODES, TEXINFO, FACES, LIGHTING (RGB), CLIPNODES (empty),
LEAFS, MARKSURFACES, EDGES, SURFEDGES, MODELS.

World model: a voxel grid of CELL-unit cells, each empty / solid / water.  Rooms, doorways, shafts,
pillars and pools are carved procedurally, then a real BSP tree is built by recursive axial splits
(largest wall first), faces are the merged empty|solid (and water) boundaries lying on each node's
plane, and every face is listed in the marksurfaces of every leaf it bounds - the invariants the
reference's R_RecursiveWorldNode (ref/soft/r_bsp.c:466) relies on.

Everything is integer / hash based: the same seed gives the same bytes on every machine.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np

CELL = 32
EMPTY, SOLID, WATER = 0, 1, 2
CONTENTS = {EMPTY: -1, SOLID: -2, WATER: -3}
MAX_FACE_CELLS = 7  # 7*32 = 224 <= 240 texels -> lightmap <= 16x16 luxels
TEX_SPECIAL = 1


def _hash_u32(*arrs):
    """Vectorised integer hash (xorshift-multiply) of integer arrays -> uint32."""
    h = np.uint64(0x9E3779B97F4A7C15)
    with np.errstate(over="ignore"):
        acc = np.zeros(np.broadcast(*arrs).shape, dtype=np.uint64) + h
        for a in arrs:
            acc = acc ^ (np.asarray(a).astype(np.int64).astype(np.uint64) + np.uint64(0x9E3779B9) + (acc << np.uint64(6)) + (acc >> np.uint64(2)))
            acc = acc * np.uint64(0xBF58476D1CE4E5B9)
            acc = acc ^ (acc >> np.uint64(31))
        return (acc & np.uint64(0xFFFFFFFF)).astype(np.uint32)


class LCG:
    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFF

    def next(self):
        self.s = (self.s * 1664525 + 1013904223) & 0xFFFFFFFF
        return self.s >> 8

    def randint(self, lo, hi):  # inclusive
        return lo + self.next() % (hi - lo + 1)

    def chance(self, num, den):
        return self.next() % den < num

    def uniform(self, lo, hi):
        return lo + (hi - lo) * (self.next() & 0xFFFF) / 65535.0


# ---------------------------------------------------------------------------------------------
# textures


def _value_noise(w, h, seed, octaves=4):
    ys, xs = np.mgrid[0:h, 0:w]
    out = np.zeros((h, w), dtype=np.float64)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        step = max(1, min(w, h) >> (o + 2))
        gx, gy = xs // step, ys // step
        fx, fy = (xs % step) / step, (ys % step) / step
        nx, ny = max(1, w // step), max(1, h // step)

        def g(ix, iy):
            return (_hash_u32(ix % nx, iy % ny, np.int64(seed + o * 7919)) & 0xFFFF) / 65535.0

        sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        v = (g(gx, gy) * (1 - sx) + g(gx + 1, gy) * sx) * (1 - sy) + (g(gx, gy + 1) * (1 - sx) + g(gx + 1, gy + 1) * sx) * sy
        out += v * amp
        tot += amp
        amp *= 0.5
    return out / tot


def make_miptex(name: str, w: int, h: int, seed: int, kind: str = "wall") -> bytes:
    """One embedded miptex: mip_t header, 4 mip levels of palette indices, 256-colour palette."""
    n = _value_noise(w, h, seed)
    ys, xs = np.mgrid[0:h, 0:w]
    rng = LCG(seed * 2654435761)
    base = np.array([rng.randint(60, 255), rng.randint(60, 255), rng.randint(60, 255)], dtype=np.float64)
    alt = np.array([rng.randint(0, 160), rng.randint(0, 160), rng.randint(0, 160)], dtype=np.float64)
    if kind == "wall":
        bw, bh = max(8, w // 4), max(4, h // 8)
        row = ys // bh
        mortar = ((ys % bh) < 1) | (((xs + (row % 2) * (bw // 2)) % bw) < 1)
        n = n * 0.7 + 0.3 * ((_hash_u32((xs + (row % 2) * (bw // 2)) // bw, row, np.int64(seed)) & 0xFF) / 255.0)
        n = np.where(mortar, 0.05, n)
    elif kind == "floor":
        tile = max(8, w // 4)
        n = n * 0.8 + 0.2 * (((xs // tile) + (ys // tile)) % 2)
    elif kind == "water":
        n = 0.5 + 0.5 * np.sin(xs * 2 * np.pi / w * 2 + n * 6) * np.cos(ys * 2 * np.pi / h * 2)
        base, alt = np.array([60.0, 140.0, 235.0]), np.array([10.0, 40.0, 110.0])
    elif kind == "sky":
        base, alt = np.array([150.0, 190.0, 255.0]), np.array([90.0, 120.0, 220.0])
    idx = np.clip((n * 255.0).astype(np.int64), 0, 254).astype(np.uint8)  # 255 is reserved ('{' alpha index)
    t = (np.arange(256, dtype=np.float64) / 255.0)[:, None]
    pal = np.clip(alt[None, :] * (1 - t) + base[None, :] * t, 0, 255).astype(np.uint8)
    mips = [idx]
    for lvl in range(1, 4):
        mips.append(idx[:: 1 << lvl, :: 1 << lvl].copy())
    hdr_size = 40
    offs, o = [], hdr_size
    for m in mips:
        offs.append(o)
        o += m.size
    blob = struct.pack("<16sII4I", name.encode("ascii"), w, h, *offs)
    for m in mips:
        blob += m.tobytes()
    blob += struct.pack("<H", 256) + pal.tobytes()
    blob += b"\0" * ((-len(blob)) % 4)
    return blob


# ---------------------------------------------------------------------------------------------
# world description


@dataclass
class MapSpec:
    rooms: tuple = (6, 6, 2)  # rooms along x, y, z
    room_cells: tuple = (8, 8, 6)  # pitch in cells incl. one wall cell
    seed: int = 0xB200
    door_prob: tuple = (3, 4)
    shaft_prob: tuple = (1, 5)
    pillar_prob: tuple = (1, 2)
    water_prob: tuple = (0, 1)  # pools in ground-floor rooms (C5)
    sky_prob: tuple = (0, 1)
    extra_styles_prob: tuple = (0, 1)  # faces that get light styles 1..3 in addition to 0 (C3)
    ntextures: int = 16
    name: str = "synth"
    novis: bool = False  # no visibility lump: every leaf's visofs is -1
    brush_models: int = 0  # box-shaped brush submodels "*1".."*n" (doors / crates for brush entities), centred on their own origin


@dataclass
class GeneratedMap:
    bsp: bytes
    spec: MapSpec
    stats: dict
    cameras: np.ndarray = field(default_factory=lambda: np.zeros((0, 6), np.float32))  # x y z pitch yaw roll
    grid: np.ndarray | None = None
    origin: tuple = (0, 0, 0)
    extras: dict | None = None  # what the generator placed besides the world: brush submodels, sprite / model spots


def _carve(spec: MapSpec):
    rx, ry, rz = spec.rooms
    cx, cy, cz = spec.room_cells
    nx, ny, nz = rx * cx + 1, ry * cy + 1, rz * cz + 1
    g = np.full((nx, ny, nz), SOLID, dtype=np.int8)
    zone = np.full((nx, ny, nz), -1, dtype=np.int32)
    rng = LCG(spec.seed)
    nzones = 0
    room_zone = {}
    for k in range(rz):
        for j in range(ry):
            for i in range(rx):
                x0, y0, z0 = i * cx + 1, j * cy + 1, k * cz + 1
                g[x0 : x0 + cx - 1, y0 : y0 + cy - 1, z0 : z0 + cz - 1] = EMPTY
                zone[x0 : x0 + cx - 1, y0 : y0 + cy - 1, z0 : z0 + cz - 1] = nzones
                room_zone[(i, j, k)] = nzones
                nzones += 1
    # doorways along x and y
    for k in range(rz):
        for j in range(ry):
            for i in range(rx):
                z0 = k * cz + 1
                if i + 1 < rx and rng.chance(*spec.door_prob):
                    wx = (i + 1) * cx
                    dy = j * cy + 1 + rng.randint(1, cy - 4)
                    dh = rng.randint(3, cz - 2)
                    g[wx, dy : dy + 2, z0 : z0 + dh] = EMPTY
                    zone[wx, dy : dy + 2, z0 : z0 + dh] = nzones
                    nzones += 1
                if j + 1 < ry and rng.chance(*spec.door_prob):
                    wy = (j + 1) * cy
                    dx = i * cx + 1 + rng.randint(1, cx - 4)
                    dh = rng.randint(3, cz - 2)
                    g[dx : dx + 2, wy, z0 : z0 + dh] = EMPTY
                    zone[dx : dx + 2, wy, z0 : z0 + dh] = nzones
                    nzones += 1
                if k + 1 < rz and rng.chance(*spec.shaft_prob):
                    wz = (k + 1) * cz
                    sx = i * cx + 1 + rng.randint(0, cx - 4)
                    sy = j * cy + 1 + rng.randint(0, cy - 4)
                    g[sx : sx + 2, sy : sy + 2, wz] = EMPTY
                    zone[sx : sx + 2, sy : sy + 2, wz] = nzones
                    nzones += 1
    # pillars and pools (pools: sunken one cell into the slab below ground-floor rooms is impossible at
    # the bottom slab, so a pool is a raised rim-less layer of water on the room floor, one cell deep)
    for k in range(rz):
        for j in range(ry):
            for i in range(rx):
                x0, y0, z0 = i * cx + 1, j * cy + 1, k * cz + 1
                if rng.chance(*spec.pillar_prob):
                    px, py = x0 + rng.randint(1, cx - 4), y0 + rng.randint(1, cy - 4)
                    if (g[px, py, z0 - 1] == SOLID) and (g[px, py, z0 + cz - 1] == SOLID):
                        g[px, py, z0 : z0 + cz - 1] = SOLID
                        zone[px, py, z0 : z0 + cz - 1] = -1
                if k == 0 and spec.water_prob[0] and rng.chance(*spec.water_prob):
                    sub = g[x0 : x0 + cx - 1, y0 : y0 + cy - 1, z0]
                    sub[sub == EMPTY] = WATER
    return g, zone, nzones, room_zone


# ---------------------------------------------------------------------------------------------
# BSP construction


class _Builder:
    def __init__(self, spec: MapSpec, g, zone, nzones):
        self.spec, self.g, self.zone, self.nzones = spec, g, zone, nzones
        self.dims = g.shape
        # world origin so that coordinates are centred (keeps int16 node bounds small)
        self.org = tuple(-(d * CELL) // 2 // CELL * CELL for d in self.dims)
        self.planes, self.plane_idx = [], {}
        self.nodes, self.leafs = [], []
        self.faces = []  # dict(plane, side, rect(axis,c,u0,u1,v0,v1), tex, node)
        self.leaf_faces = {}
        self.leafs.append(dict(contents=-2, box=None, faces=[]))  # shared solid leaf 0
        self.rng = LCG(spec.seed ^ 0x5F3759DF)

    def plane(self, axis, c):
        key = (axis, c)
        if key not in self.plane_idx:
            self.plane_idx[key] = len(self.planes)
            self.planes.append(key)
        return self.plane_idx[key]

    def world(self, axis, c):
        return self.org[axis] + c * CELL

    def build(self, box):
        x0, x1, y0, y1, z0, z1 = box
        sub = self.g[x0:x1, y0:y1, z0:z1]
        first = sub.flat[0]
        if (sub == first).all():
            if first == SOLID:
                return -1  # leaf 0
            self.leafs.append(dict(contents=CONTENTS[int(first)], box=box, faces=[]))
            return -(len(self.leafs) - 1) - 1
        # choose the split: plane with most content transitions, tie -> closest to the middle
        best = None
        for axis in range(3):
            n = sub.shape[axis]
            if n < 2:
                continue
            a = np.moveaxis(sub, axis, 0)
            diff = (a[1:] != a[:-1]).reshape(n - 1, -1).sum(axis=1)
            if diff.max() == 0:
                continue
            mid = (n - 1) / 2.0
            for i in np.flatnonzero(diff == diff.max()):
                score = (int(diff[i]), -abs(i + 1 - mid - 0.5), -axis)
                if best is None or score > best[0]:
                    best = (score, axis, int(i) + 1)
        _, axis, off = best
        lo = (x0, y0, z0)[axis]
        c = lo + off
        ni = len(self.nodes)
        node = dict(plane=self.plane(axis, c), box=box, children=[0, 0], firstface=0, numfaces=0, axis=axis, c=c)
        self.nodes.append(node)
        self._emit_faces(ni, node)
        fb = list(box)
        bb = list(box)
        fb[axis * 2] = c  # front: coord >= c
        bb[axis * 2 + 1] = c
        node["children"][0] = self.build(tuple(fb))
        node["children"][1] = self.build(tuple(bb))
        return ni

    def _emit_faces(self, ni, node):
        axis, c = node["axis"], node["c"]
        x0, x1, y0, y1, z0, z1 = node["box"]
        lo = [x0, y0, z0]
        hi = [x1, y1, z1]
        ua, va = [a for a in range(3) if a != axis]
        sl_f = [slice(lo[a], hi[a]) for a in range(3)]
        sl_b = list(sl_f)
        sl_f[axis] = c
        sl_b[axis] = c - 1
        front = self.g[tuple(sl_f)]
        back = self.g[tuple(sl_b)]
        node["firstface"] = len(self.faces)
        # (mask, side, kind): side 0 = normal +axis visible from the front half-space
        jobs = [
            ((front == EMPTY) & (back == SOLID), 0, "solid"),
            ((back == EMPTY) & (front == SOLID), 1, "solid"),
            ((front == WATER) & (back == SOLID), 0, "solid"),
            ((back == WATER) & (front == SOLID), 1, "solid"),
            ((front == EMPTY) & (back == WATER), 0, "water"),
            ((front == EMPTY) & (back == WATER), 1, "water"),
            ((back == EMPTY) & (front == WATER), 1, "water"),
            ((back == EMPTY) & (front == WATER), 0, "water"),
        ]
        for mask, side, kind in jobs:
            if not mask.any():
                continue
            for (u0, u1, v0, v1) in _rectangles(mask):
                self.faces.append(
                    dict(plane=node["plane"], side=side, axis=axis, c=c, u=(lo[ua] + u0, lo[ua] + u1), v=(lo[va] + v0, lo[va] + v1), kind=kind, node=ni)
                )
        node["numfaces"] = len(self.faces) - node["firstface"]

    # ---- leaf <-> face adjacency (marksurfaces)
    def mark(self):
        for fi, f in enumerate(self.faces):
            node = self.nodes[f["node"]]
            # the leaf that sees the face lies on the side its normal points to
            child = node["children"][0 if f["side"] == 0 else 1]
            self._mark_down(child, fi, f)

    def _mark_down(self, child, fi, f):
        if child < 0:
            li = -1 - child
            if li == 0:
                return
            leaf = self.leafs[li]
            if self._touches(leaf["box"], f):
                leaf["faces"].append(fi)
            return
        node = self.nodes[child]
        if not self._touches(node["box"], f):
            return
        self._mark_down(node["children"][0], fi, f)
        self._mark_down(node["children"][1], fi, f)

    @staticmethod
    def _touches(box, f):
        axis, c = f["axis"], f["c"]
        lo, hi = box[axis * 2], box[axis * 2 + 1]
        if f["side"] == 0:
            if lo != c:
                return False
        elif hi != c:
            return False
        ua, va = [a for a in range(3) if a != axis]
        return min(box[ua * 2 + 1], f["u"][1]) > max(box[ua * 2], f["u"][0]) and min(box[va * 2 + 1], f["v"][1]) > max(box[va * 2], f["v"][0])


def _rectangles(mask):
    """Greedy decomposition of a boolean 2-D mask into rectangles of at most MAX_FACE_CELLS a side."""
    m = mask.copy()
    out = []
    nu, nv = m.shape
    for u in range(nu):
        v = 0
        while v < nv:
            if not m[u, v]:
                v += 1
                continue
            v1 = v
            while v1 < nv and m[u, v1] and v1 - v < MAX_FACE_CELLS:
                v1 += 1
            u1 = u + 1
            while u1 < nu and u1 - u < MAX_FACE_CELLS and m[u1, v:v1].all():
                u1 += 1
            m[u:u1, v:v1] = False
            out.append((u, u1, v, v1))
            v = v1
    return out


# texture axes per plane axis, Quake style (s, t) with unit scale
_TEXVECS = {
    0: ((0.0, 1.0, 0.0, 0.0), (0.0, 0.0, -1.0, 0.0)),
    1: ((1.0, 0.0, 0.0, 0.0), (0.0, 0.0, -1.0, 0.0)),
    2: ((1.0, 0.0, 0.0, 0.0), (0.0, -1.0, 0.0, 0.0)),
}


def _face_vertices(b: _Builder, f):
    """Clockwise (as seen from the visible side) corner list in world units."""
    axis, c = f["axis"], f["c"]
    n = np.zeros(3)
    n[axis] = 1.0 if f["side"] == 0 else -1.0
    ua, va = [a for a in range(3) if a != axis]
    up = np.zeros(3)
    up[va] = 1.0
    right = np.cross(-n, up)  # viewer looks along -n
    ra = int(np.flatnonzero(right)[0])
    assert ra == ua
    rs = right[ra]
    u0, u1 = b.world(ua, f["u"][0]), b.world(ua, f["u"][1])
    v0, v1 = b.world(va, f["v"][0]), b.world(va, f["v"][1])
    rl, rr = (u0, u1) if rs > 0 else (u1, u0)
    pc = b.world(axis, c)

    def P(r, t):
        p = [0, 0, 0]
        p[axis], p[ua], p[va] = pc, r, t
        return tuple(p)

    return [P(rl, v1), P(rr, v1), P(rr, v0), P(rl, v0)]


def generate(spec: MapSpec = MapSpec(), ncameras: int = 256) -> GeneratedMap:
    g, zone, nzones, room_zone = _carve(spec)
    b = _Builder(spec, g, zone, nzones)
    root = b.build((0, g.shape[0], 0, g.shape[1], 0, g.shape[2]))
    assert root == 0
    b.mark()
    rng = LCG(spec.seed ^ 0xC0FFEE)

    # ---- textures
    ntex = spec.ntextures
    tex_blobs, tex_names, tex_kind = [], [], []
    sizes = [64, 128, 256]
    for i in range(ntex):
        kind = ("wall", "floor", "wall", "floor")[i % 4]
        w = sizes[(i // 2) % 3]
        h = sizes[(i // 3) % 3]
        name = f"{kind}{i:02d}"
        tex_names.append(name)
        tex_kind.append(kind)
        tex_blobs.append(make_miptex(name, w, h, spec.seed + i * 101, kind))
    water_tex = len(tex_names)
    tex_names.append("!water0")
    tex_kind.append("water")
    tex_blobs.append(make_miptex("!water0", 64, 64, spec.seed + 7777, "water"))
    sky_tex = len(tex_names)
    tex_names.append("sky")
    tex_kind.append("sky")
    tex_blobs.append(make_miptex("sky", 64, 64, spec.seed + 8888, "sky"))
    wall_texs = [i for i, k in enumerate(tex_kind) if k == "wall"]
    floor_texs = [i for i, k in enumerate(tex_kind) if k == "floor"]

    texinfo, texinfo_idx = [], {}

    def get_texinfo(axis, miptex, flags):
        key = (axis, miptex, flags)
        if key not in texinfo_idx:
            texinfo_idx[key] = len(texinfo)
            texinfo.append((*_TEXVECS[axis][0], *_TEXVECS[axis][1], miptex, flags))
        return texinfo_idx[key]

    # ---- geometry: shared vertices and edges
    verts, vert_idx = [], {}
    edges, edge_dir = [(0, 0)], {}
    surfedges = []

    def vid(p):
        if p not in vert_idx:
            vert_idx[p] = len(verts)
            verts.append(p)
        return vert_idx[p]

    def eid(a, c):
        if (c, a) in edge_dir and (a, c) not in edge_dir:
            e = edge_dir[(c, a)]
            edge_dir[(a, c)] = -e
            return -e
        if (a, c) in edge_dir:  # direction already used by another face: a fresh edge keeps <=2 users
            edges.append((a, c))
            return len(edges) - 1
        edges.append((a, c))
        edge_dir[(a, c)] = len(edges) - 1
        return len(edges) - 1

    cx, cy, cz = spec.room_cells
    light_blob = bytearray()
    face_recs = []
    lights = []  # one point light per room, world units
    for (i, j, k), z in room_zone.items():
        lights.append(
            (
                b.world(0, i * cx + 1) + (cx - 1) * CELL / 2,
                b.world(1, j * cy + 1) + (cy - 1) * CELL / 2,
                b.world(2, k * cz + 1) + (cz - 1) * CELL * 0.8,
                100 + (int(_hash_u32(np.int64(i), np.int64(j), np.int64(k + spec.seed))) % 156),
                int(_hash_u32(np.int64(i * 3 + 1), np.int64(j * 5 + 2), np.int64(k * 7 + spec.seed))) % 3,
            )
        )
    lights_a = np.array(lights, dtype=np.float64)

    for fi, f in enumerate(b.faces):
        axis = f["axis"]
        corners = _face_vertices(b, f)
        room = (f["u"][0] // 8 * 131 + f["v"][0] // 8 * 71 + f["c"] // 8 * 31 + axis) & 0xFFFF
        flags = 0
        styles = [0, 255, 255, 255]
        if f["kind"] == "water":
            miptex, flags = water_tex, TEX_SPECIAL
        elif axis == 2:
            miptex = floor_texs[room % len(floor_texs)]
            if f["side"] == 1 and spec.sky_prob[0] and f["c"] == g.shape[2] - 1 and (room * 2654435761 >> 7) % spec.sky_prob[1] < spec.sky_prob[0]:
                miptex, flags = sky_tex, TEX_SPECIAL
        else:
            miptex = wall_texs[room % len(wall_texs)]
        ti = get_texinfo(axis, miptex, flags)
        first = len(surfedges)
        ids = [vid(p) for p in corners]
        for q in range(4):
            surfedges.append(eid(ids[q], ids[(q + 1) % 4]))
        lightofs = -1
        if not flags & TEX_SPECIAL:
            sv, tv = _TEXVECS[axis]
            P = np.array(corners, dtype=np.float64)
            s = P @ np.array(sv[:3]) + sv[3]
            t = P @ np.array(tv[:3]) + tv[3]
            smin, tmin = int(np.floor(s.min() / 16)), int(np.floor(t.min() / 16))
            smax_, tmax_ = int(np.ceil(s.max() / 16)), int(np.ceil(t.max() / 16))
            sw, th = smax_ - smin + 1, tmax_ - tmin + 1
            if spec.extra_styles_prob[0] and (int(_hash_u32(np.int64(fi), np.int64(spec.seed))) % spec.extra_styles_prob[1]) < spec.extra_styles_prob[0]:
                nst = 2 + int(_hash_u32(np.int64(fi), np.int64(spec.seed + 1))) % 3
                styles = [0] + [1 + int(_hash_u32(np.int64(fi), np.int64(q))) % 11 for q in range(1, nst)] + [255] * (4 - nst)
                # a face must not list the same style twice
                seen, st2 = set(), []
                for sst in styles:
                    if sst != 255 and sst in seen:
                        continue
                    seen.add(sst)
                    st2.append(sst)
                styles = (st2 + [255] * 4)[:4]
            nstyles = sum(1 for sst in styles if sst != 255)
            # luxel world positions
            ss = (smin + np.arange(sw)) * 16.0
            tt = (tmin + np.arange(th)) * 16.0
            S, T = np.meshgrid(ss, tt)  # [th, sw]
            W = np.zeros((th, sw, 3))
            W[..., axis] = b.world(axis, f["c"])
            sa = int(np.flatnonzero(np.array(sv[:3]))[0])
            ta = int(np.flatnonzero(np.array(tv[:3]))[0])
            W[..., sa] = S * sv[sa]
            W[..., ta] = T * tv[ta]
            # lights further than the cut-off radius add exactly 0; on big maps only the ones that can reach the
            # face are summed (small maps keep the full sum: their bytes are pinned by committed goldens)
            la = lights_a
            if len(lights_a) > 100:
                wmin, wmax = W.reshape(-1, 3).min(0), W.reshape(-1, 3).max(0)
                gap = np.maximum(np.maximum(wmin[None, :] - lights_a[:, :3], lights_a[:, :3] - wmax[None, :]), 0.0)
                la = lights_a[(gap**2).sum(-1) < (6 * CELL * 1.5) ** 2]
            d2 = ((W[:, :, None, :] - la[None, None, :, :3]) ** 2).sum(-1)
            contrib = la[None, None, :, 3] / (1.0 + d2 / (140.0**2))
            near = d2 < (6 * CELL * 1.5) ** 2
            contrib = contrib * near
            lightofs = len(light_blob)
            for si in range(nstyles):
                rgb = np.zeros((th, sw, 3))
                for ch in range(3):
                    tint = 1.0 - 0.35 * (la[None, None, :, 4] == ch)
                    rgb[..., ch] = (contrib * tint).sum(-1)
                noise = (_hash_u32(S.astype(np.int64), T.astype(np.int64), np.int64(fi * 4 + si)) & 0x1F).astype(np.float64)
                amb = 18.0 if si == 0 else 0.0
                scale = 1.0 if si == 0 else 0.6
                rgb = np.clip(rgb * scale + amb + noise[..., None] * (1.0 if si == 0 else 0.5), 0, 255).astype(np.uint8)
                light_blob += rgb.tobytes()
        face_recs.append((f["plane"], f["side"], first, 4, ti, styles, lightofs))

    # ---- brush submodels: boxes around their own origin (the entity's origin places them), six outward faces each
    nworldfaces = len(face_recs)
    extra_planes = []  # (axis, dist) in model space, after the world's planes
    submodels = []  # (mins, maxs, firstface, numfaces)
    for mi in range(spec.brush_models):
        half = (16 + 8 * (int(_hash_u32(np.int64(mi), np.int64(spec.seed + 11))) % 4),
                16 + 8 * (int(_hash_u32(np.int64(mi), np.int64(spec.seed + 12))) % 4),
                16 + 16 * (int(_hash_u32(np.int64(mi), np.int64(spec.seed + 13))) % 3))
        first_face = len(face_recs)
        for axis in range(3):
            for side in (0, 1):
                nsign = 1.0 if side == 0 else -1.0
                pc = nsign * half[axis]
                ua, va = [a for a in range(3) if a != axis]
                n = np.zeros(3)
                n[axis] = nsign
                up = np.zeros(3)
                up[va] = 1.0
                right = np.cross(-n, up)
                rs = right[ua]
                rl, rr = (-half[ua], half[ua]) if rs > 0 else (half[ua], -half[ua])

                def P(r, t, axis=axis, ua=ua, va=va, pc=pc):
                    q = [0.0, 0.0, 0.0]
                    q[axis], q[ua], q[va] = pc, r, t
                    return ("m", mi, *q)  # keyed apart from the world's vertices

                corners = [P(rl, half[va]), P(rr, half[va]), P(rr, -half[va]), P(rl, -half[va])]
                miptex = (wall_texs if axis != 2 else floor_texs)[(mi + axis) % len(wall_texs if axis != 2 else floor_texs)]
                ti = get_texinfo(axis, miptex, 0)
                first = len(surfedges)
                ids = [vid(c) for c in corners]
                for q in range(4):
                    surfedges.append(eid(ids[q], ids[(q + 1) % 4]))
                sv, tv = _TEXVECS[axis]
                Pm = np.array([c[2:] for c in corners], dtype=np.float64)
                sc_ = Pm @ np.array(sv[:3]) + sv[3]
                tc_ = Pm @ np.array(tv[:3]) + tv[3]
                smin, tmin = int(np.floor(sc_.min() / 16)), int(np.floor(tc_.min() / 16))
                sw, th = int(np.ceil(sc_.max() / 16)) - smin + 1, int(np.ceil(tc_.max() / 16)) - tmin + 1
                S, T = np.meshgrid(np.arange(sw), np.arange(th))
                base = 90 + 25 * axis + (40 if side == 0 else 0)
                noise = (_hash_u32(S.astype(np.int64), T.astype(np.int64), np.int64(mi * 16 + axis * 2 + side + 9000)) & 0x1F).astype(np.float64)
                lum = np.clip(base + 6.0 * S + 4.0 * T + noise, 0, 255)
                rgb = np.stack([lum, np.clip(lum * 0.9 + 10, 0, 255), np.clip(lum * 0.8 + (mi % 3) * 20, 0, 255)], -1).astype(np.uint8)
                lightofs = len(light_blob)
                light_blob += rgb.tobytes()
                extra_planes.append((axis, float(pc)))
                face_recs.append((len(b.planes) + len(extra_planes) - 1, side, first, 4, ti, [0, 255, 255, 255], lightofs))
        submodels.append((tuple(-float(h) for h in half), tuple(float(h) for h in half), first_face, 6))

    # ---- marksurfaces / leafs
    marks = []
    leaf_recs = []
    for li, leaf in enumerate(b.leafs):
        if li == 0:
            leaf_recs.append((-2, -1, (0, 0, 0), (0, 0, 0), 0, 0))
            continue
        box = leaf["box"]
        mins = tuple(b.world(a, box[a * 2]) for a in range(3))
        maxs = tuple(b.world(a, box[a * 2 + 1]) for a in range(3))
        leaf_recs.append((leaf["contents"], 0, mins, maxs, len(marks), len(leaf["faces"])))
        marks.extend(leaf["faces"])
    nleafs = len(b.leafs)
    nvis = nleafs - 1

    # ---- PVS: zones, monotone flood (a sight line never reverses direction along any axis)
    vis_rows = _compute_pvs(b, g, zone, nzones, nvis)
    vis_blob = bytearray()
    visofs = [-1] * nleafs
    for li in range(1, nleafs):
        if spec.novis:
            break
        visofs[li] = len(vis_blob)
        vis_blob += _rle(vis_rows[li - 1])

    # ---- serialise
    L = {}
    ent = '{\n"classname" "worldspawn"\n"message" "%s"\n}\n' % spec.name
    L[0] = ent.encode("ascii") + b"\0"
    L[1] = b"".join(struct.pack("<4fi", *[(1.0 if a == ax else 0.0) for a in range(3)], float(b.world(ax, c)), ax) for (ax, c) in b.planes)
    L[1] += b"".join(struct.pack("<4fi", *[(1.0 if a == ax else 0.0) for a in range(3)], d, ax) for (ax, d) in extra_planes)
    nt = len(tex_blobs)
    hdr = struct.pack("<i", nt)
    o = 4 + 4 * nt
    offs = []
    for blob in tex_blobs:
        offs.append(o)
        o += len(blob)
    L[2] = hdr + struct.pack(f"<{nt}i", *offs) + b"".join(tex_blobs)
    L[3] = b"".join(struct.pack("<3f", *map(float, p[2:] if p[0] == "m" else p)) for p in verts)
    L[4] = bytes(vis_blob)
    node_blob = bytearray()
    for n in b.nodes:
        box = n["box"]
        mins = [b.world(a, box[a * 2]) for a in range(3)]
        maxs = [b.world(a, box[a * 2 + 1]) for a in range(3)]
        ch = [c if c >= 0 else c for c in n["children"]]  # negative = -(leaf+1) already
        node_blob += struct.pack("<i2h3h3hHH", n["plane"], ch[0], ch[1], *mins, *maxs, n["firstface"], n["numfaces"])
    # brush submodels: a chain of six nodes each, one per face plane - outside of a plane ends in a leaf, inside goes on to the
    # next plane (the engine walks a submodel's own tree from headnode[0]: decal placement, R_DecalNode)
    sub_heads = []
    nnodes_total = len(b.nodes)
    for (mn, mx, ff, nf) in submodels:
        sub_heads.append(nnodes_total)
        for q in range(nf):
            planenum, side = face_recs[ff + q][0], face_recs[ff + q][1]
            nxt = nnodes_total + 1 if q + 1 < nf else -1  # -1 = -(leaf 0 + 1)
            ch = (-1, nxt) if side == 0 else (nxt, -1)
            node_blob += struct.pack("<i2h3h3hHH", planenum, ch[0], ch[1], *[int(v) for v in mn], *[int(v) for v in mx], ff + q, 1)
            nnodes_total += 1
    L[5] = bytes(node_blob)
    L[6] = b"".join(struct.pack("<8fihh", *ti[:8], ti[8], ti[9], -1) for ti in texinfo)
    L[7] = b"".join(struct.pack("<HhihH4Bi", p, s, fe, ne, ti, *st, lo) for (p, s, fe, ne, ti, st, lo) in face_recs)
    L[8] = bytes(light_blob)
    L[9] = b""
    L[10] = b"".join(
        struct.pack("<ii3h3hHH4B", c, visofs[i] if i else -1, *mn, *mx, fm, nm, 0, 0, 0, 0) for i, (c, _, mn, mx, fm, nm) in enumerate(leaf_recs)
    )
    L[11] = struct.pack(f"<{len(marks)}H", *marks)
    L[12] = b"".join(struct.pack("<HH", a, c) for (a, c) in edges)
    L[13] = struct.pack(f"<{len(surfedges)}i", *surfedges)
    wmins = [b.world(a, 0) for a in range(3)]
    wmaxs = [b.world(a, g.shape[a]) for a in range(3)]
    L[14] = struct.pack("<9f4i3i", *map(float, wmins), *map(float, wmaxs), 0.0, 0.0, 0.0, 0, 0, 0, 0, nvis, 0, nworldfaces)
    for k, (mn, mx, ff, nf) in enumerate(submodels):
        L[14] += struct.pack("<9f4i3i", *mn, *mx, 0.0, 0.0, 0.0, sub_heads[k], 0, 0, 0, 0, ff, nf)
    assert len(verts) < 65536 and len(face_recs) < 65536 and len(marks) < 65536 and nnodes_total < 32768 and nleafs < 32768
    out = bytearray(struct.pack("<i", 30) + b"\0" * (8 * 15))
    for i in range(15):
        while len(out) % 4:
            out += b"\0"
        struct.pack_into("<ii", out, 4 + 8 * i, len(out), len(L[i]))
        out += L[i]

    cams = _cameras(b.org, g, spec, ncameras)
    stats = dict(
        faces=len(face_recs), nodes=len(b.nodes), leafs=nleafs, planes=len(b.planes), verts=len(verts), edges=len(edges),
        marksurfaces=len(marks), textures=nt, lightbytes=len(light_blob), visbytes_row=(nvis + 7) // 8, vislump=len(vis_blob),
        bsp_bytes=len(out),
    )
    extras = {"brush_models": [{"modelindex": 2 + i, "mins": mn, "maxs": mx} for i, (mn, mx, _, _) in enumerate(submodels)]}
    return GeneratedMap(bytes(out), spec, stats, cams, g, b.org, extras)


def _rle(row: np.ndarray) -> bytes:
    """Zero-run RLE of a PVS row (Mod_DecompressPVS's inverse): non-zero bytes verbatim, runs of zeros as {0, count <= 255}."""
    row = np.asarray(row, dtype=np.uint8)
    n = len(row)
    if not n:
        return b""
    nz = (row != 0).astype(np.int8)
    cuts = np.concatenate(([0], np.flatnonzero(np.diff(nz)) + 1, [n]))
    out = bytearray()
    raw = row.tobytes()
    for a, c in zip(cuts[:-1].tolist(), cuts[1:].tolist()):
        if raw[a]:
            out += raw[a:c]
        else:
            full, rem = divmod(c - a, 255)
            out += b"\x00\xff" * full
            if rem:
                out += bytes((0, rem))
    return bytes(out)


def _compute_pvs(b: _Builder, g, zone, nzones, nvis):
    """Leaf-to-leaf PVS rows (bit i = leaf i+1) from a conservative zone flood."""
    # zone adjacency with crossing direction
    adj = [dict() for _ in range(nzones)]  # zone -> {(axis, dir): set(zone)}
    for axis in range(3):
        a = np.moveaxis(zone, axis, 0)
        za, zb = a[:-1], a[1:]
        m = (za >= 0) & (zb >= 0) & (za != zb)
        pa, pb = za[m], zb[m]
        for z0, z1 in set(zip(pa.tolist(), pb.tolist())):
            adj[z0].setdefault((axis, 1), set()).add(z1)
            adj[z1].setdefault((axis, -1), set()).add(z0)
    # a path is valid iff along every axis all its moves go the same way <=> all moves lie in one of the 8 octants:
    # per octant the reachable set is plain graph reachability over that octant's directed edges
    from scipy.sparse import csr_matrix
    from scipy.sparse.csgraph import shortest_path

    zvis = np.eye(nzones, dtype=bool)
    for octant in range(8):
        rows, cols = [], []
        for axis in range(3):
            d = 1 if (octant >> axis) & 1 else -1
            for z0 in range(nzones):
                nb = adj[z0].get((axis, d))
                if nb:
                    rows.extend([z0] * len(nb))
                    cols.extend(nb)
        if not rows:
            continue
        a = csr_matrix((np.ones(len(rows), dtype=np.int8), (rows, cols)), shape=(nzones, nzones))
        for lo in range(0, nzones, 1024):  # blocks of sources bound the dense distance matrix
            idx = np.arange(lo, min(nzones, lo + 1024))
            dist = shortest_path(a, method="D", directed=True, unweighted=True, indices=idx)
            zvis[idx] |= np.isfinite(dist)
    # leaf -> zones
    nleafs = len(b.leafs)
    lr, lc = [], []
    for li in range(1, nleafs):
        x0, x1, y0, y1, z0, z1 = b.leafs[li]["box"]
        zs = np.unique(zone[x0:x1, y0:y1, z0:z1])
        zs = zs[zs >= 0]
        lr.extend([li] * len(zs))
        lc.extend(zs.tolist())
    lz = csr_matrix((np.ones(len(lr), dtype=np.float32), (lr, lc)), shape=(nleafs, nzones))
    nbytes = (nvis + 7) // 8
    rows = []
    lzt = lz.T.tocsr()
    for lo in range(1, nleafs, 2048):
        hi = min(nleafs, lo + 2048)
        seen_z = lz[lo:hi] @ zvis.astype(np.float32)  # [leaf, zone]: zones the leaf's zones see
        vis = (csr_matrix(seen_z > 0, dtype=np.float32) @ lzt).toarray() > 0  # [leaf, leaf]
        vis[np.arange(hi - lo), np.arange(lo, hi)] = True
        packed = np.packbits(vis[:, 1:], axis=1, bitorder="little")
        rows.extend(packed[i, :nbytes] for i in range(hi - lo))
    return rows


def _cameras(org, g, spec: MapSpec, n: int):
    """Eye points inside empty space, >= 1 cell from any wall sideways; yaw U[0,360), pitch U[-30,30]."""
    rng = LCG(spec.seed ^ 0xCA3E7A)
    cams = []
    nx, ny, nz = g.shape
    cx, cy, cz = spec.room_cells
    guard = 0
    while len(cams) < n and guard < n * 200:
        guard += 1
        x, y = rng.randint(1, nx - 2), rng.randint(1, ny - 2)
        k = rng.randint(0, spec.rooms[2] - 1)
        z = k * cz + 2  # second cell above the floor ~ eye height 48..64
        if not (g[x - 1 : x + 2, y - 1 : y + 2, z] == EMPTY).all():
            continue
        px = org[0] + x * CELL + rng.uniform(4, CELL - 4)
        py = org[1] + y * CELL + rng.uniform(4, CELL - 4)
        pz = org[2] + z * CELL + rng.uniform(4, CELL - 4)
        cams.append((px, py, pz, rng.uniform(-30, 30), rng.uniform(0, 360), 0.0))
    return np.array(cams, dtype=np.float32)


# ---------------------------------------------------------------------------------------------
# named configurations (BASELINE.json configs)

def spec_small(seed=0xB200, **kw) -> MapSpec:
    """Map S: 6x6x2 rooms, ~2k faces (C1, C2, C5)."""
    return MapSpec(rooms=(6, 6, 2), seed=seed, name="synth_s", **kw)


def spec_tiny(seed=0xB200, **kw) -> MapSpec:
    """2x2x1 rooms for unit tests."""
    return MapSpec(rooms=(2, 2, 1), seed=seed, name="synth_t", **kw)


def spec_medium(seed=0xB203, **kw) -> MapSpec:
    """Map M: 12x12x3 rooms, ~12.5k faces, extra light styles on 25% of faces: the C3 mechanism (1080p, animated
    light styles, cache rebuilds) at a size the generator finishes in ~20 s (Map L takes many minutes to generate)."""
    kw.setdefault("extra_styles_prob", (1, 4))
    return MapSpec(rooms=(12, 12, 3), seed=seed, name="synth_m", **kw)


def spec_large(seed=0xB203, **kw) -> MapSpec:
    """Map L: ~30k faces, extra light styles 1..11 on 25% of faces (C3).  16x16x4 rooms: the 24x24x4 grid SURVEY.md
    pencilled in gives ~66k faces / ~100k vertices with this room generator, past BSP30's 16-bit vertex and
    marksurface indices (common/bspfile.h:214-222); BASELINE.json states the face count, and this grid has it."""
    kw.setdefault("extra_styles_prob", (1, 4))
    return MapSpec(rooms=(16, 16, 4), seed=seed, name="synth_l", **kw)


def generate_cached(spec: MapSpec = MapSpec(), ncameras: int = 256, cache_dir: str | None = None) -> GeneratedMap:
    """generate() behind a disk cache keyed by the spec and this file's own bytes (big maps take a minute to build;
    the cache directory is git-ignored and travels with the tree to the GPU box)."""
    import hashlib
    import os
    import pickle

    cache_dir = cache_dir or os.path.join(os.path.dirname(os.path.abspath(__file__)), "_mapcache")
    with open(os.path.abspath(__file__), "rb") as f:
        key = hashlib.sha256(f.read() + repr(spec).encode()).hexdigest()[:20]
    path = os.path.join(cache_dir, f"{spec.name}_{key}.pkl")
    if os.path.exists(path):
        try:
            with open(path, "rb") as f:
                bsp, stats, grid, origin, extras = pickle.load(f)
            gm = GeneratedMap(bsp, spec, stats, _cameras(origin, grid, spec, ncameras), grid, origin)
            gm.extras = extras
            return gm
        except Exception:
            pass
    gm = generate(spec, ncameras)
    if len(gm.bsp) > (1 << 20):  # small maps generate in a second or two
        try:
            os.makedirs(cache_dir, exist_ok=True)
            tmp = path + f".{os.getpid()}.tmp"
            with open(tmp, "wb") as f:
                pickle.dump((gm.bsp, gm.stats, gm.grid, gm.origin, getattr(gm, "extras", None)), f)
            os.replace(tmp, path)
        except OSError:
            pass
    return gm


if __name__ == "__main__":
    import sys
    import time

    which = sys.argv[1] if len(sys.argv) > 1 else "small"
    t0 = time.time()
    m = generate({"small": spec_small, "tiny": spec_tiny, "large": spec_large}[which]())
    print(m.stats, f"{time.time() - t0:.1f}s", "cameras", len(m.cameras))
    if len(sys.argv) > 2:
        open(sys.argv[2], "wb").write(m.bsp)
