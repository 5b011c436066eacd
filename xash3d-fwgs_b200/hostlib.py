"""ctypes binding of libxashhost.so, the headless engine stand-in (harness/xh_host.h at the repository root: test and bench infrastructure).

One `Host` == one engine instance == one private copy of the library image (the renderer ABI has
no context pointer and renderers keep process-global state), so two renderers - e.g. the oracle
ref_soft and ref_cuda - can live in one process.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
HOST_LIB_PATH = os.path.join(os.path.dirname(_HERE), "harness", "libxashhost.so")
REF_CUDA_PATH = os.path.join(_HERE, "libref_cuda.so")
REF_SOFT_PATH = os.path.join(_ROOT, "oracle", "_ref", "libref_soft.so")

RF_DRAW_WORLD = 1


def calc_fov_y(fov_x: float, width: float, height: float) -> float:
    """engine/client/cl_view.c V_CalcFov: fov_y from fov_x and the aspect ratio."""
    import math

    if fov_x < 1 or fov_x > 179:
        fov_x = 90.0
    x = width / math.tan(fov_x / 360.0 * math.pi)
    return math.atan(height / x) * 360.0 / math.pi


class Host:
    _copies = 0
    _loaded_refs: set = set()

    def __init__(self, ref_path: str, width: int, height: int, cvars: dict | None = None, sw_format=None):
        if not os.path.exists(HOST_LIB_PATH):
            raise RuntimeError(f"{HOST_LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        if not os.path.exists(ref_path):
            raise RuntimeError(f"renderer library {ref_path} missing (no CPU fallback exists): build it first")
        # private image of the host library for this instance
        Host._copies += 1
        self._tmp = tempfile.NamedTemporaryFile(prefix=f"xashhost{Host._copies}_", suffix=".so", delete=False)
        self._tmp.close()
        shutil.copyfile(HOST_LIB_PATH, self._tmp.name)
        self.lib = C.CDLL(self._tmp.name)
        os.unlink(self._tmp.name)
        L = self.lib
        L.xh_init.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.xh_last_error.restype = C.c_char_p
        L.xh_cvar_set.argtypes = [C.c_char_p, C.c_char_p]
        L.xh_cvar_get.argtypes = [C.c_char_p]
        L.xh_cvar_get.restype = C.c_float
        L.xh_load_map.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p]
        L.xh_set_time.argtypes = [C.c_double, C.c_double]
        L.xh_set_lightstyle.argtypes = [C.c_int, C.c_char_p, C.c_float]
        L.xh_render_view.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.xh_ref_symbol.argtypes = [C.c_char_p]
        L.xh_ref_symbol.restype = C.c_void_p
        L.xh_world_model.restype = C.c_void_p
        L.xh_world_info.argtypes = [C.c_void_p, C.c_int]
        L.xh_point_in_leaf.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.xh_surface_info.argtypes = [C.c_int, C.c_void_p, C.c_int]
        L.xh_gamma_table.argtypes = [C.c_int, C.c_void_p]
        L.xh_gamma_table.restype = C.c_void_p
        L.xh_build_gamma.argtypes = [C.c_float] * 4
        L.xh_callback_count.restype = C.c_uint64
        L.xh_soft_cache_block.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.xh_soft_cache_block.restype = C.c_void_p
        L.xh_add_brush_entity.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        self.width, self.height = width, height
        self.ref_path = ref_path
        # (bpp, rmask, gmask, bmask) of the software surface SW_CreateBuffer hands out; default RGB565
        self.sw_format = tuple(sw_format) if sw_format else (2, 0xF800, 0x07E0, 0x001F)
        L.xh_set_sw_format.argtypes = [C.c_uint] * 4
        L.xh_set_sw_format(*self.sw_format)
        L.xh_sw_buffer.restype = C.c_void_p
        L.xh_sw_buffer.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.xh_soft_blit_views.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        for k, v in (cvars or {}).items():
            L.xh_cvar_set(k.encode(), str(v).encode())
        # renderers keep process-global state: a second instance of the same plugin in this process gets a
        # private image of it too (the first one is loaded in place, from the tree)
        load_path, tmp_ref = ref_path, None
        if os.path.realpath(ref_path) in Host._loaded_refs:
            t = tempfile.NamedTemporaryFile(prefix=f"xashref{Host._copies}_", suffix=".so", delete=False)
            t.close()
            shutil.copyfile(ref_path, t.name)
            load_path = tmp_ref = t.name
        Host._loaded_refs.add(os.path.realpath(ref_path))
        rc = L.xh_init(load_path.encode(), width, height)
        if tmp_ref:
            os.unlink(tmp_ref)
        if rc != 0:
            raise RuntimeError(f"xh_init({ref_path}) failed: {rc} {L.xh_last_error().decode()}")
        self._map = None

    # -- engine-side state
    def cvar_set(self, name, value):
        self.lib.xh_cvar_set(name.encode(), str(value).encode())

    def cvar_get(self, name) -> float:
        return float(self.lib.xh_cvar_get(name.encode()))

    def load_map(self, bsp: bytes, name="synth"):
        self._map = C.create_string_buffer(bsp, len(bsp))  # the engine keeps the file image alive
        rc = self.lib.xh_load_map(self._map, len(bsp), name.encode())
        if rc != 0:
            raise RuntimeError(f"xh_load_map failed: {rc}")

    def world_info(self) -> dict:
        a = (C.c_int * 16)()
        n = self.lib.xh_world_info(a, 16)
        keys = ["numsurfaces", "numnodes", "numleafs", "numedges", "numvertexes", "numsurfedges", "nummarksurfaces", "numplanes", "numtextures", "numsubmodels", "visbytes"]
        return dict(zip(keys, list(a)[:n]))

    def surface_info(self, i) -> dict:
        a = (C.c_int * 16)()
        n = self.lib.xh_surface_info(i, a, 16)
        keys = ["flags", "extents0", "extents1", "lightextents0", "lightextents1", "style0", "style1", "style2", "style3", "texnum", "texturemins0", "texturemins1", "numedges", "has_samples"]
        return dict(zip(keys, list(a)[:n]))

    def set_time(self, t, oldt=None):
        self.lib.xh_set_time(t, t - 1.0 / 60 if oldt is None else oldt)

    def set_lightstyle(self, i, pattern: str, time=0.0):
        self.lib.xh_set_lightstyle(i, pattern.encode(), time)

    def run_lightstyles(self):
        self.lib.xh_run_lightstyles()

    def gamma_table(self, which) -> np.ndarray:
        n = C.c_int()
        p = self.lib.xh_gamma_table(which, C.byref(n))
        dt = np.uint8 if which == 0 else np.uint16
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8 if which == 0 else C.c_uint16)), (n.value,)).astype(dt).copy()

    def build_gamma(self, gamma, brightness, texgamma, lightgamma):
        self.lib.xh_build_gamma(gamma, brightness, texgamma, lightgamma)

    def symbol(self, name) -> int:
        return self.lib.xh_ref_symbol(name.encode()) or 0

    # -- frames
    def render_view(self, origin, angles, fov_x=90.0, fov_y=None, w=None, h=None, flags=RF_DRAW_WORLD):
        w = w or self.width
        h = h or self.height
        if fov_y is None:
            fov_y = calc_fov_y(fov_x, w, h)
        o = (C.c_float * 3)(*map(float, origin))
        a = (C.c_float * 3)(*map(float, angles))
        self.lib.xh_render_view(o, a, fov_x, fov_y, 0, 0, w, h, flags)

    def begin_frame(self):
        self.lib.xh_begin_frame()

    def end_frame(self):
        self.lib.xh_end_frame()

    # -- ref_soft (oracle) buffer probes: vid.buffer / d_pzbuffer are plain exported globals
    def soft_buffers(self):
        vid = self.symbol("vid")
        zb = self.symbol("d_pzbuffer")
        assert vid and zb, "not a ref_soft image"
        cptr = C.cast(vid, C.POINTER(C.c_void_p))[0]
        zptr = C.cast(zb, C.POINTER(C.c_void_p))[0]
        n = self.width * self.height
        col = np.ctypeslib.as_array(C.cast(cptr, C.POINTER(C.c_uint16)), (self.height, self.width))
        z = np.ctypeslib.as_array(C.cast(zptr, C.POINTER(C.c_int16)), (self.height, self.width))
        return col, z

    def soft_reset_view_cache(self):
        """Make every view independent of the previous one: forget the cached view cluster
        (R_MarkLeaves early-out, ref/soft/r_main.c:1023) so the fat PVS is rebuilt per view."""
        p = self.symbol("r_oldviewcluster")
        C.cast(p, C.POINTER(C.c_int))[0] = -2

    def soft_cache_block(self, surf, mip):
        w, h = C.c_int(), C.c_int()
        p = self.lib.xh_soft_cache_block(surf, mip, C.byref(w), C.byref(h))
        if not p:
            return None
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint16)), (h.value, w.value)).copy()


def pixels_to_rgb(buf: np.ndarray) -> np.ndarray:
    """ref_soft pixel_t (major RGB332 << 8 | minor GBRGBRGB, ref/soft/r_local.h:84) -> RGB888 for viewing."""
    p = buf.astype(np.uint32)
    m, j = p >> 8, p & 0xFF
    bit = lambda v, b: (v >> b) & 1
    r = (((m >> 5) << 2) & 31) | (bit(j, 5) << 1) | bit(j, 2)
    g = (((m >> 2) << 3) & 63) | (bit(j, 7) << 2) | (bit(j, 4) << 1) | bit(j, 1)
    b = ((m << 3) & 31) | (bit(j, 6) << 2) | (bit(j, 3) << 1) | bit(j, 0)
    return np.stack([r * 255 // 31, g * 255 // 63, b * 255 // 31], -1).astype(np.uint8)


def screen_map(sw_format) -> np.ndarray:
    """Numpy restatement of R_BuildScreenMap (ref/soft/r_glblit.c:347-407), test infrastructure: the 65536-entry table
    pixel_t -> display pixel for sw_format = (bpp, rmask, gmask, bmask).  Entry (i << 8 | j) = major(i) | minor(j)."""
    _, rmask, gmask, bmask = sw_format
    shift = lambda m: (m & -m).bit_length() - 1 if m else 0
    mult = lambda m: 1 << bin(m).count("1")
    i = np.arange(256, dtype=np.uint64)
    bit = lambda v, b: (v >> np.uint64(b)) & np.uint64(1)
    scale = lambda c, m, d: ((c * np.uint64(mult(m))) // np.uint64(d)) << np.uint64(shift(m))
    r = ((i >> np.uint64(5)) << np.uint64(2)) & np.uint64(31)
    g = ((i >> np.uint64(2)) << np.uint64(3)) & np.uint64(63)
    b = (i << np.uint64(3)) & np.uint64(31)
    major = scale(r, rmask, 31) | scale(g, gmask, 63) | scale(b, bmask, 31)
    r = (bit(i, 5) << np.uint64(1)) | bit(i, 2)
    g = (bit(i, 7) << np.uint64(2)) | (bit(i, 4) << np.uint64(1)) | bit(i, 1)
    b = (bit(i, 6) << np.uint64(2)) | (bit(i, 3) << np.uint64(1)) | bit(i, 0)
    minor = scale(r, rmask, 31) | scale(g, gmask, 63) | scale(b, bmask, 31)
    return (major[:, None] | minor[None, :]).reshape(65536).astype(np.uint32)
