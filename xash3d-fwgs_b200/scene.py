"""Synthetic scenes + one `Renderer` front end for any renderer plugin behind the headless host.

`Renderer(REF_CUDA_PATH, ...)` drives libref_cuda.so (the product: batched views on the GPU through
the RefCuda_RenderViews C-ABI entry, include/ref_cuda.h).  `Renderer(REF_SOFT_PATH, ...)` drives the
unmodified reference ref_soft view by view through GL_RenderFrame (engine/ref_api.h:544) - that is the
parity oracle and the CPU baseline; only tests, smoke() and bench.py's reference legs construct it.
"""
from __future__ import annotations

import ctypes as C
import functools
import importlib

import math

import numpy as np

from .hostlib import Host, REF_CUDA_PATH, REF_SOFT_PATH, RF_DRAW_WORLD, calc_fov_y

mapgen = importlib.import_module(__package__ + ".tools.mapgen")


class RefViewpass(C.Structure):
    """common/ref_params.h:97 ref_viewpass_t"""

    _fields_ = [
        ("viewport", C.c_int * 4),
        ("vieworigin", C.c_float * 3),
        ("viewangles", C.c_float * 3),
        ("viewentity", C.c_int),
        ("fov_x", C.c_float),
        ("fov_y", C.c_float),
        ("flags", C.c_int),
    ]


assert C.sizeof(RefViewpass) == 56

CTR_NAMES = ["views", "pixels", "cands", "surfs", "edges", "spans", "cache_blocks", "cache_texels",
             "overflow_edges", "overflow_events", "overflow_stack", "overflow_cands", "serial_lines", "sorted_lines", "unstaged_bands",
             "max_cands", "max_edges", "max_band_events", "overflow_cache", "tris", "tri_pixels", "wide_lines", "overflow_fix", "fixed_faces", "tri_written", "big_tris", "bmodel_surfs", "bmodel_ties", "overflow_bmodel"]


@functools.lru_cache(maxsize=8)
def make_map(which: str = "small", ncameras: int = 256, **kw):
    """Named synthetic worlds of BASELINE.json: tiny (unit tests), small (Map S: C1/C2/C5), large (Map L: C3)."""
    spec = {"tiny": mapgen.spec_tiny, "small": mapgen.spec_small, "medium": mapgen.spec_medium, "large": mapgen.spec_large}[which](**kw)
    return mapgen.generate_cached(spec, ncameras=ncameras)


class Renderer:
    def __init__(self, ref_path: str, width: int, height: int, gmap=None, cvars: dict | None = None, sw_format=None):
        self.host = Host(ref_path, width, height, cvars, sw_format)
        # the batch extension (include/ref_cuda.h) is what makes a plugin drivable in batches; ref_soft has none
        self.is_cuda = bool(self.host.symbol("RefCuda_RenderViews"))
        self.width, self.height = width, height
        L = self.host.lib
        L.xh_soft_render_views.restype = C.c_double
        L.xh_soft_render_views.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        if self.is_cuda:
            sym = self.host.symbol
            self._render = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int)(sym("RefCuda_RenderViews"))
            self._render_async = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int)(sym("RefCuda_RenderViewsAsync"))
            self._wait = C.CFUNCTYPE(C.c_int)(sym("RefCuda_Wait"))
            self._bench = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float))(sym("RefCuda_BenchResident"))
            self._readcache = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int))(sym("RefCuda_ReadSurfaceCache"))
            self._counters = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)(sym("RefCuda_GetCounters"))
            self._launches = C.CFUNCTYPE(C.c_ulonglong)(sym("RefCuda_LaunchCount"))
            self._devname = C.CFUNCTYPE(C.c_char_p)(sym("RefCuda_DeviceName"))
            self._h2d = C.CFUNCTYPE(C.c_ulonglong)(sym("RefCuda_BatchH2DBytes"))
            self._blit = C.CFUNCTYPE(C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.POINTER(C.c_float))(sym("RefCuda_BlitViews"))
        if gmap is not None:
            self.load_map(gmap)

    def load_map(self, gmap):
        self.map = gmap
        self.host.load_map(gmap.bsp, gmap.spec.name)
        self.host.run_lightstyles()

    # -- time / light styles (engine side)
    def set_time(self, t):
        self.host.set_time(t)

    def set_particles(self, parts):
        """the client's active particles: rows of {x, y, z, palette index, die time, ptype_t}; drawn by CL_DrawParticles in every
        following batch (needs a texture named '*particle': load_pic)"""
        a = np.ascontiguousarray(np.asarray(parts, dtype=np.float32).reshape(-1, 6))
        L = self.host.lib
        L.xh_set_particles.argtypes = [C.c_void_p, C.c_int]
        return L.xh_set_particles(a.ctypes.data, len(a))

    def set_sight_wall(self, axis, dist=0.0):
        """the wall the engine's sight line (EV_VisTraceLine, glow sprites) stops at: coordinate[axis] = dist; axis < 0: none"""
        L = self.host.lib
        L.xh_set_sight_wall.argtypes = [C.c_int, C.c_float]
        L.xh_set_sight_wall.restype = None
        L.xh_set_sight_wall(int(axis), float(dist))

    def world_to_screen(self, p):
        """the renderer's WorldToScreen slot for the view rendered last -> (behind, x, y)"""
        L = self.host.lib
        L.xh_world_to_screen.argtypes = [C.c_void_p, C.c_void_p]
        pp = (C.c_float * 3)(*[float(x) for x in p])
        out = (C.c_float * 3)()
        b = L.xh_world_to_screen(pp, out)
        return (int(b), float(out[0]), float(out[1]))

    def set_tracers(self, rows, dot_model):
        """the client's active tracers: rows of {org[3], vel[3], ramp, die, colour index, ptype_t, alpha}; drawn by
        CL_DrawTracers in every following view with the sprite model dot_model as REF_DOT_SPRITE"""
        a = np.ascontiguousarray(np.asarray(rows, dtype=np.float32).reshape(-1, 11))
        L = self.host.lib
        L.xh_set_dot_sprite.argtypes = [C.c_int]
        L.xh_set_dot_sprite.restype = None
        L.xh_set_dot_sprite(int(dot_model))
        L.xh_set_tracers.argtypes = [C.c_void_p, C.c_int]
        return L.xh_set_tracers(a.ctypes.data, len(a))

    def get_tracers(self, n):
        """the tracer list as the renderer's last CL_DrawTracers call left it: rows of {org[3], vel[3], alpha}"""
        out = np.zeros((n, 7), np.float32)
        L = self.host.lib
        L.xh_get_tracers.argtypes = [C.c_void_p]
        got = L.xh_get_tracers(out.ctypes.data)
        return out[:got]

    def decal_shoot_entity(self, texture, entity, pos, flags=0, scale=1.0):
        """R_DecalShoot on a brush entity (its number as add_entity returned it), pos in world space"""
        L = self.host.lib
        L.xh_decal_shoot_entity.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_float]
        L.xh_decal_shoot_entity.restype = None
        pp = (C.c_float * 3)(*[float(x) for x in pos])
        L.xh_decal_shoot_entity(int(texture), int(entity), pp, int(flags), float(scale))

    def decal_shoot(self, texture, pos, flags=0, scale=1.0):
        """the renderer's R_DecalShoot slot on the world model: a decal texture (load_pic) stamped at pos"""
        L = self.host.lib
        L.xh_decal_shoot.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_float]
        L.xh_decal_shoot.restype = None
        pp = (C.c_float * 3)(*[float(x) for x in pos])
        L.xh_decal_shoot(int(texture), pp, int(flags), float(scale))

    def decal_list(self, max_entries=4096):
        """the renderer's R_CreateDecalList slot -> rows of {position[3], entity, depth, flags, scale, plane normal[3], name checksum}"""
        out = np.zeros((max_entries, 11), np.float32)
        L = self.host.lib
        L.xh_decal_list.argtypes = [C.c_void_p, C.c_int]
        n = L.xh_decal_list(out.ctypes.data, max_entries)
        return out[:n]

    def remove_decals(self, texture=0):
        L = self.host.lib
        L.xh_remove_decals.argtypes = [C.c_int]
        L.xh_remove_decals.restype = None
        L.xh_remove_decals(int(texture))

    def clear_decals(self):
        L = self.host.lib
        L.xh_clear_decals.restype = None
        L.xh_clear_decals()

    def light_point(self, p):
        """the renderer's R_LightPoint slot: light of the world on the ground below p -> (r, g, b)"""
        L = self.host.lib
        L.xh_light_point.argtypes = [C.c_void_p, C.c_void_p]
        pp = (C.c_float * 3)(*[float(x) for x in p])
        out = (C.c_int * 4)()
        L.xh_light_point(pp, out)
        return tuple(out[:3])

    def gamma_changed(self, gamma=2.5, brightness=0.0, texgamma=2.0, lightgamma=2.5):
        """the engine's V_CheckGamma: the four gamma tables rebuilt in place, then R_GammaChanged( false )"""
        L = self.host.lib
        L.xh_gamma_changed.argtypes = [C.c_float] * 4
        L.xh_gamma_changed(gamma, brightness, texgamma, lightgamma)

    def set_lightstyle(self, i, pattern, time=0.0):
        self.host.set_lightstyle(i, pattern, time)

    def set_dlight(self, i, origin, radius, color=(255, 255, 255), die=1e9, minlight=0.0):
        """Entry i of the engine's dynamic light table (dlight_t); radius 0 switches it off."""
        L = self.host.lib
        L.xh_set_dlight.argtypes = [C.c_int, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float]
        o = (C.c_float * 3)(*[float(x) for x in origin])
        if L.xh_set_dlight(i, o, float(radius), int(color[0]), int(color[1]), int(color[2]), float(die), float(minlight)) != 0:
            raise ValueError("dlight index out of range")

    def run_lightstyles(self):
        self.host.run_lightstyles()

    # -- entities (engine side: the client adds them to the renderer's lists before it asks for views)
    def create_sprite(self, name, rgba, sprtype=2, texformat=0, texflags=0):
        """A one-frame sprite model from an RGBA uint8 image [h, w, 4]; sprtype: SPR_* orientation (2 = SPR_FWD_PARALLEL),
        texformat: 0 normal, 1 additive, 2 indexalpha, 3 alphatest; texflags: TF_* of the frame texture.  -> model index"""
        rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
        h, w = rgba.shape[:2]
        L = self.host.lib
        L.xh_create_sprite.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        mi = L.xh_create_sprite(name.encode(), w, h, rgba.ctypes.data, sprtype, texformat, texflags)
        if mi <= 0:
            raise RuntimeError("xh_create_sprite failed")
        return mi

    def load_studio(self, name, mdl: bytes):
        """A studio model from an MDL v10 file image (tools/mdlgen.py writes synthetic ones).  -> model index"""
        L = self.host.lib
        L.xh_load_studio.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
        mi = L.xh_load_studio(name.encode(), mdl, len(mdl))
        if mi <= 0:
            raise RuntimeError("xh_load_studio failed")
        return mi

    def add_entity(self, modelindex, origin, angles=(0, 0, 0), rendermode=0, renderamt=255, rendercolor=(0, 0, 0), renderfx=0,
                   frame=0, sequence=0, body=0, skin=0, scale=1.0, framerate=1.0, animtime=0.0, movetype=0, aiment=0):
        ip = (C.c_int * 13)(modelindex, rendermode, renderamt, *[int(c) for c in rendercolor], renderfx, frame, sequence, body, skin, movetype, aiment)
        fp = (C.c_float * 9)(*[float(x) for x in origin], *[float(x) for x in angles], float(scale), float(framerate), float(animtime))
        L = self.host.lib
        L.xh_add_entity.argtypes = [C.c_void_p, C.c_void_p]
        idx = L.xh_add_entity(ip, fp)
        if idx < 0:
            raise RuntimeError(f"xh_add_entity failed ({idx})")
        return idx

    def clear_entities(self):
        self.host.lib.xh_clear_scene()

    # -- views
    def viewpasses(self, cams, fov_x=90.0, flags=RF_DRAW_WORLD, viewport=None):
        cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
        x, y, w, h = viewport or (0, 0, self.width, self.height)
        arr = (RefViewpass * len(cams))()
        fov_y = calc_fov_y(fov_x, w, h)
        for i, c in enumerate(cams):
            v = arr[i]
            v.viewport[:] = [x, y, w, h]
            v.vieworigin[:] = [float(c[0]), float(c[1]), float(c[2])]
            v.viewangles[:] = [float(c[3]), float(c[4]), float(c[5])]
            v.viewentity = 1
            v.fov_x, v.fov_y, v.flags = fov_x, fov_y, flags
        return arr

    def render_views(self, cams, fov_x=90.0, clear=True, out=None, flags=RF_DRAW_WORLD, viewport=None, want_z=True):
        """-> (colour[n,h,w] uint16 pixel_t, z[n,h,w] int16).  `out` = preallocated (colour, z) host arrays
        (e.g. pinned) to receive the device->host copies."""
        cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
        n = len(cams)
        if out is None:
            color = np.empty((n, self.height, self.width), np.uint16)
            z = np.empty((n, self.height, self.width), np.int16) if want_z else None
        else:
            color, z = out
        zp = z.ctypes.data if z is not None else None
        if self.is_cuda:
            vp = self.viewpasses(cams, fov_x, flags, viewport)
            rc = self._render(C.addressof(vp), n, color.ctypes.data if color is not None else None, zp, 1 if clear else 0)
            if rc != 0:
                raise RuntimeError("RefCuda_RenderViews failed: " + self.host.lib.xh_last_error().decode(errors="replace"))
            self.last_seconds = None
        else:
            x, y, w, h = viewport or (0, 0, self.width, self.height)
            cc = np.ascontiguousarray(cams)
            self.last_seconds = self.host.lib.xh_soft_render_views(cc.ctypes.data, n, fov_x, calc_fov_y(fov_x, w, h), x, y, w, h, flags,
                                                                  1 if clear else 0, color.ctypes.data, zp)
            if self.last_seconds < 0:
                raise RuntimeError("not a ref_soft image")
        return color, z

    def blit_views(self, cams, fov_x=90.0, sw_format=None, out=None):
        """Display tail (R_BlitScreen): the views in the display format `sw_format` = (bpp, rmask, gmask, bmask)
        (default: the format the host's software surface was created with) -> array[n,h,w] of uint16 / uint32.
        CUDA: the batch is rendered, stays in HBM and is converted there (RefCuda_BlitViews).  Oracle: every view goes
        through GL_RenderFrame + R_EndFrame and the software surface is copied out (its format is fixed at R_Init)."""
        cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
        n = len(cams)
        fmt = tuple(sw_format) if sw_format else self.host.sw_format
        disp = out if out is not None else np.empty((n, self.height, self.width), np.uint16 if fmt[0] == 2 else np.uint32)
        if self.is_cuda:
            self.render_resident(cams, fov_x)
            ms = C.c_float()
            if self._blit(0, n, disp.ctypes.data, self.width, fmt[0], fmt[1], fmt[2], fmt[3], C.byref(ms)) != 0:
                raise RuntimeError("RefCuda_BlitViews failed")
            self.last_blit_ms = ms.value
        else:
            if fmt != self.host.sw_format:
                raise ValueError("the oracle's display format is fixed when the renderer is created (sw_format=...)")
            cc = np.ascontiguousarray(cams)
            rc = self.host.lib.xh_soft_blit_views(cc.ctypes.data, n, fov_x, calc_fov_y(fov_x, self.width, self.height), 0, 0,
                                                  self.width, self.height, RF_DRAW_WORLD, disp.ctypes.data)
            if rc != 0:
                raise RuntimeError(f"xh_soft_blit_views failed ({rc})")
        return disp

    def frame_views(self, cams, fov_x=90.0):
        """The stock one-view-per-frame path the engine drives: R_BeginFrame, GL_RenderFrame, R_EndFrame per camera;
        -> the software surface after each frame, array[n,h,w] in the host's display format."""
        cams = np.ascontiguousarray(np.asarray(cams, dtype=np.float32).reshape(-1, 6))
        n = len(cams)
        disp = np.empty((n, self.height, self.width), np.uint16 if self.host.sw_format[0] == 2 else np.uint32)
        L = self.host.lib
        L.xh_frame_views.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
        if L.xh_frame_views(cams.ctypes.data, n, fov_x, calc_fov_y(fov_x, self.width, self.height), 0, 0, self.width, self.height,
                            RF_DRAW_WORLD, disp.ctypes.data) != 0:
            raise RuntimeError("xh_frame_views failed")
        return disp

    def load_pic(self, name, rgba, texflags=0):
        """an RGBA image (h, w, 4 uint8) as a renderer texture (HUD pictures) -> texture number"""
        rgba = np.ascontiguousarray(rgba, dtype=np.uint8)
        L = self.host.lib
        L.xh_load_pic.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        t = L.xh_load_pic(name.encode(), rgba.shape[1], rgba.shape[0], rgba.ctypes.data, texflags)
        if t <= 0:
            raise RuntimeError("xh_load_pic failed")
        return t

    def frame_views_hud(self, cams, hud, fov_x=90.0):
        """frame_views with a HUD pass after every view.  hud: rows of {kind 0 pic / 1 fill, rendermode, r, g, b, a, x, y, w, h,
        s1, t1, s2, t2, texnum}: GL_SetRenderMode + Color4ub + R_DrawStretchPic, or FillRGBA."""
        cams = np.ascontiguousarray(np.asarray(cams, dtype=np.float32).reshape(-1, 6))
        hud = np.ascontiguousarray(np.asarray(hud, dtype=np.float32).reshape(-1, 15))
        n = len(cams)
        disp = np.empty((n, self.height, self.width), np.uint16 if self.host.sw_format[0] == 2 else np.uint32)
        L = self.host.lib
        L.xh_frame_views_hud.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_void_p, C.c_int, C.c_void_p]
        if L.xh_frame_views_hud(cams.ctypes.data, n, fov_x, calc_fov_y(fov_x, self.width, self.height), 0, 0, self.width, self.height,
                                RF_DRAW_WORLD, hud.ctypes.data, len(hud), disp.ctypes.data) != 0:
            raise RuntimeError("xh_frame_views_hud failed")
        return disp

    def screenshot(self, shot_type=0):
        """VID_ScreenShot of the frame last shown by R_EndFrame -> the BGRA image FS_SaveImage received, array[h, w] uint32"""
        L = self.host.lib
        L.xh_screenshot.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        buf = np.empty(((self.height + 3) & ~3) * ((self.width + 3) & ~3), np.uint32)
        w, h = C.c_int(), C.c_int()
        n = L.xh_screenshot(shot_type, buf.ctypes.data, buf.nbytes, C.byref(w), C.byref(h))
        if n < 0:
            raise RuntimeError(f"xh_screenshot failed ({n})")
        return buf[: w.value * h.value].reshape(h.value, w.value)

    def render_resident(self, cams, fov_x=90.0, clear=True):
        """Render with the results left in HBM (no device->host copy)."""
        assert self.is_cuda
        vp = self.viewpasses(cams, fov_x)
        if self._render(C.addressof(vp), len(vp), None, None, 1 if clear else 0) != 0:
            raise RuntimeError("RefCuda_RenderViews failed")

    def render_prepared(self, vp, color_ptr, z_ptr, clear=True):
        """RefCuda_RenderViews on a prebuilt ref_viewpass_t array and raw host pointers (bench e2e leg)."""
        if self._render(C.addressof(vp), len(vp), color_ptr, z_ptr, 1 if clear else 0) != 0:
            raise RuntimeError("RefCuda_RenderViews failed")

    def render_prepared_async(self, vp, color_ptr, z_ptr, clear=True):
        """RefCuda_RenderViewsAsync: enqueued only; the host buffers are complete after wait()"""
        if self._render_async(C.addressof(vp), len(vp), color_ptr, z_ptr, 1 if clear else 0) != 0:
            raise RuntimeError("RefCuda_RenderViewsAsync failed")

    def wait(self):
        if self._wait() != 0:
            raise RuntimeError("RefCuda_Wait failed")

    def bench_resident(self, iters, rebuild_cache=False, flush_l2=False, overlap=False, serial_cache=False):
        """CUDA-event ms per iteration of the last batch re-run with inputs resident in HBM: (total, [6 stages: vis_walk,
        faces_edges, surface_cache, spans, draw, entities]).  overlap=True: the slicing RefCuda_RenderViews uses for host outputs (4 view
        slices on 2 streams): total only, the stage list is zeros."""
        tot = C.c_float()
        st = (C.c_float * 6)()
        if self._bench(iters, (1 if rebuild_cache else 0) | (2 if flush_l2 else 0) | (4 if overlap else 0) | (8 if serial_cache else 0), C.byref(tot), st) != 0:
            raise RuntimeError("RefCuda_BenchResident failed")
        return tot.value, list(st)

    # -- probes
    def cache_block(self, surf, mip):
        if not self.is_cuda:
            return self.host.soft_cache_block(surf, mip)
        w, h, valid = C.c_int(), C.c_int(), C.c_int()
        buf = np.empty(512 * 512, np.uint16)
        if self._readcache(surf, mip, buf.ctypes.data, buf.size, C.byref(w), C.byref(h), C.byref(valid)) != 0:
            return None
        if not valid.value:
            return None
        return buf[: w.value * h.value].reshape(h.value, w.value).copy()

    def counters(self) -> dict:
        a = np.zeros(32, np.uint64)
        self._counters(a.ctypes.data, 32)
        return dict(zip(CTR_NAMES, [int(x) for x in a[: len(CTR_NAMES)]]))

    def batch_h2d_bytes(self) -> int:
        """host->device bytes the last RefCuda_RenderViews call uploaded (per-view constants, light styles, entity state)"""
        return int(self._h2d())

    def launch_count(self) -> int:
        return int(self._launches())

    def device_name(self) -> str:
        return self._devname().decode()


def add_studio_models(r, cams, count=64, nbones=32, ahead=72.0, drop=56.0, mdl=None, animtime=0.5):
    """BASELINE config C4's entities: `count` instances of the synthetic skinned model (tools/mdlgen.py: `nbones`
    bones, ~1500 vertices, two meshes) standing `ahead` units in front of the first `count` cameras, each at its own
    yaw and animation phase.  The same call on a ref_cuda and a ref_soft Renderer gives both the same scene."""
    from .tools import mdlgen
    if mdl is None:
        mdl = mdlgen.generate(nbones=nbones)
    mi = r.load_studio("models/synth.mdl", mdl)
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        org = (c[0] + ahead * math.cos(yaw), c[1] + ahead * math.sin(yaw), c[2] - drop)
        r.add_entity(mi, org, angles=(0.0, 37.0 * k, 0.0), frame=(k * 3) % 30, framerate=1.0, animtime=animtime + 0.01 * (k % 16))
    return mi


def add_brush_entities(r, m, cams, count=16, ahead=88.0, drop=24.0, rotate=True, rendermode=0, renderamt=255):
    """brush entities (doors / crates): instances of the map's box submodels (make_map(..., brush_models=n)) `ahead` units
    in front of the first `count` cameras; every second one turned about the vertical axis, every fifth tilted as well."""
    models = m.extras["brush_models"]
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        org = (c[0] + ahead * math.cos(yaw), c[1] + ahead * math.sin(yaw), c[2] - drop)
        ang = (0.0, 0.0, 0.0)
        if rotate and k % 2:
            ang = (12.0 * (k % 5 == 0), 25.0 + 13.0 * k, 8.0 * (k % 5 == 0))
        r.add_entity(models[k % len(models)]["modelindex"], org, angles=ang, rendermode=rendermode, renderamt=renderamt)


def add_abutting_boxes(r, m, cams):
    """brush entities placed to make coincident edges: every box stands exactly on the floor (its bottom face in the floor's
    plane), and a second box of the same submodel touches it side by side (shared face plane, collinear top and bottom edges)"""
    models = m.extras["brush_models"]
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    for k, c in enumerate(cams):
        md = models[k % len(models)]
        yaw = np.radians(float(c[4]))
        floor = float(c[2]) - 56.0
        org = np.array([round(c[0] + 96.0 * np.cos(yaw)), round(c[1] + 96.0 * np.sin(yaw)), floor - md["mins"][2]])
        r.add_entity(md["modelindex"], tuple(org))
        r.add_entity(md["modelindex"], (org[0] + (md["maxs"][0] - md["mins"][0]), org[1], org[2]))
        if k % 3 == 0:  # a third one on top: coincident horizontal faces
            r.add_entity(md["modelindex"], (org[0], org[1], org[2] + (md["maxs"][2] - md["mins"][2])))


def add_mixed_brush_entities(r, m, cams):
    """opaque, blended (kRenderTransTexture), additive and alpha-tested brush entities in front of the first 26 cameras"""
    add_brush_entities(r, m, cams, 8, rotate=True)
    add_brush_entities(r, m, cams[8:], 6, rotate=True, rendermode=2, renderamt=140)
    add_brush_entities(r, m, cams[14:], 6, rotate=False, rendermode=5, renderamt=200)
    add_brush_entities(r, m, cams[20:], 6, rotate=True, rendermode=4, renderamt=255)


def add_particles(r, cams, count=256, ahead=60.0, seed=5):
    """`count` particles scattered in front of the cameras (palette colours, some fading out): CL_DrawParticles' input"""
    rng = np.random.default_rng(seed)
    d = np.zeros((8, 8, 4), np.uint8)
    yy, xx = np.mgrid[0:8, 0:8]
    dot = np.clip(255 - 60 * np.hypot(xx - 3.5, yy - 3.5), 0, 255).astype(np.uint8)
    d[..., 0] = d[..., 1] = d[..., 2] = dot
    d[..., 3] = 255
    r.load_pic("*particle", d)
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    rows = []
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        off = rng.uniform(-24, 24, 3)
        dist = ahead + 10.0 * (k // len(cams))
        rows.append([c[0] + dist * math.cos(yaw) + off[0], c[1] + dist * math.sin(yaw) + off[1], c[2] + off[2] * 0.5,
                     int(rng.integers(40, 256)), 1.0 + [10.0, 0.03, 0.05][k % 3], [0, 0, 6][k % 3]])  # pt_static, pt_static fading, pt_fire
    return r.set_particles(rows)


def add_glow_sprites(r, cams, count=24, seed=21):
    """kRenderGlow sprites (lamp coronas) at several distances in front of the cameras, a little right of and above the line
    of sight (ref_soft draws a glow only where R_WorldToScreen's x and y are positive), some with kRenderFxNoDissipation"""
    rng = np.random.default_rng(seed)
    mod = r.create_sprite("spr_glow", sprite_image(31, 32, 32), sprtype=2, texformat=1)
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        dist = [40.0, 90.0, 160.0, 260.0][k % 4] + float(rng.uniform(0, 20))
        side = float(rng.uniform(0.05, 0.5)) * dist
        org = (c[0] + dist * math.cos(yaw) + side * math.sin(yaw), c[1] + dist * math.sin(yaw) - side * math.cos(yaw),
               c[2] + float(rng.uniform(0.02, 0.35)) * dist)
        r.add_entity(mod, org, rendermode=3, renderamt=int(rng.integers(120, 256)), renderfx=14 if k % 5 == 0 else 0,
                     rendercolor=[(255, 220, 160), (160, 200, 255), (255, 255, 255)][k % 3], scale=float(rng.uniform(0.5, 1.5)))
    return count


def add_tracers(r, cams, count=64, ahead=70.0, seed=9, moving=False):
    """`count` tracers in front of the cameras: streaks in random directions, every colour and width of CL_DrawTracers' tables.
    moving=False keeps to the particle types the function leaves alone (it is called once per view by ref_soft and once per
    batch by ref/cuda); moving=True adds pt_grav / pt_slowgrav, whose velocity and alpha it rewrites."""
    rng = np.random.default_rng(seed)
    d = np.zeros((16, 16, 4), np.uint8)
    yy, xx = np.mgrid[0:16, 0:16]
    d[..., 0] = d[..., 1] = d[..., 2] = np.clip(255 - 28 * np.hypot(xx - 7.5, yy - 7.5), 0, 255).astype(np.uint8)
    d[..., 1] = (d[..., 1] * (0.5 + 0.5 * rng.random((16, 16)))).astype(np.uint8)
    d[..., 3] = 255
    dot = r.create_sprite("spr_dot", d, sprtype=2, texformat=1)
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    types = [0, 3, 5, 6, 9] + ([1, 2] if moving else [])  # pt_static, pt_fire, pt_explode2, pt_blob, ... (+ pt_grav, pt_slowgrav)
    rows = []
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        off = rng.uniform(-30, 30, 3)
        dist = ahead + 12.0 * (k // len(cams))
        vel = rng.uniform(-1, 1, 3)
        vel *= rng.uniform(200, 900) / np.linalg.norm(vel)
        rows.append([c[0] + dist * math.cos(yaw) + off[0], c[1] + dist * math.sin(yaw) + off[1], c[2] + off[2] * 0.4,
                     vel[0], vel[1], vel[2], rng.uniform(0.5, 1.5), 1.0 + [10.0, 0.04, 0.07][k % 3],
                     int(rng.integers(-1, 13)), types[k % len(types)], int(rng.integers(60, 256))])
    r.set_tracers(rows, dot)
    return len(rows)


def sprite_image(seed, w=64, h=64):
    """a synthetic RGBA sprite frame: a soft disc with noise, transparent outside"""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    d = np.hypot(x - w / 2, y - h / 2) / (w / 2)
    img = np.zeros((h, w, 4), np.uint8)
    img[..., 0] = np.clip(255 * (1 - d) + rng.integers(0, 40, (h, w)), 0, 255)
    img[..., 1] = np.clip(200 * (1 - d * d), 0, 255)
    img[..., 2] = rng.integers(0, 255, (h, w))
    img[..., 3] = np.where(d < 0.9, 255, 0)
    return img


def add_sprites(r, cams, count=32, ahead=96.0):
    """BASELINE config C5's sprites: `count` translucent sprites (additive, alpha-tested texture and colour-blended, all
    parallel-oriented) `ahead` units in front of the first `count` cameras; sorted per view by the renderer."""
    TF_HAS_ALPHA = 1 << 16
    mods = [r.create_sprite("spr_c5_add", sprite_image(11), sprtype=2, texformat=1),
            r.create_sprite("spr_c5_alpha", sprite_image(12), sprtype=2, texformat=3, texflags=TF_HAS_ALPHA),
            r.create_sprite("spr_c5_blend", sprite_image(13, 32, 48), sprtype=2, texformat=0)]
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    for k in range(count):
        c = cams[k % len(cams)]
        yaw = math.radians(float(c[4]))
        org = (c[0] + ahead * math.cos(yaw), c[1] + ahead * math.sin(yaw), c[2] - 8.0 + 4.0 * (k % 5))
        r.add_entity(mods[k % 3], org, rendermode=[5, 4, 2][k % 3], renderamt=[255, 255, 160][k % 3],
                     rendercolor=[(255, 200, 120), (255, 255, 255), (90, 200, 255)][k % 3], scale=[1.0, 0.75, 1.5][k % 3])


def compare_frames(a_col, a_z, b_col, b_z):
    """Fraction of identical pixels (colour, z) and whether every differing pixel sits within one pixel of
    a place where the reference image changes surface (a span boundary), per BASELINE.json's tolerance."""
    same_c = float((a_col == b_col).mean())
    same_z = float((a_z == b_z).mean()) if a_z is not None and b_z is not None else 1.0
    return same_c, same_z


def pixel_point(cam, z, x, y, fov_x=90.0):
    """the world point a pixel of a frame shows: cam = {origin, pitch, yaw, roll = 0}, z = that frame's 1/z buffer"""
    h, w = z.shape
    depth = 32768.0 / max(int(z[y, x]), 1)  # the z buffer holds 0x8000 / depth along the view axis
    p, yw = np.radians(cam[3]), np.radians(cam[4])
    fwd = np.array([np.cos(p) * np.cos(yw), np.cos(p) * np.sin(yw), -np.sin(p)])
    right = np.array([np.sin(yw), -np.cos(yw), 0.0])
    up = np.array([np.sin(p) * np.cos(yw), np.sin(p) * np.sin(yw), np.cos(p)])
    scale = (w / 2.0) / math.tan(math.radians(fov_x) / 2.0)
    d = fwd + right * ((x - (w / 2.0 - 0.5)) / scale) - up * ((y - (h / 2.0 - 0.5)) / scale)
    return np.asarray(cam[:3], dtype=np.float64) + d * depth


def shoot_decals(r, cams, z, seed=5, scale=1.0, opaque=False):
    """R_DecalShoot at the point the middle pixel of every view in cams looks at (z: 1/z frames of cams): a decal picture with
    an alpha channel (or an opaque one), on the world model.  Returns the points."""
    rng = np.random.default_rng(seed)
    TF_HAS_ALPHA = 1 << 16
    size = 32 if not opaque else 16
    img = np.zeros((size, size, 4), np.uint8)
    img[..., 0] = 255
    img[..., 1] = rng.integers(0, 255, (size, size))
    img[..., 2] = 40 * opaque
    img[..., 3] = 255 if opaque else rng.integers(0, 255, (size, size))
    if not opaque:
        img[8:24, 8:24, 3] = 255
    tex = r.load_pic("{decal%d" % seed, img, 0 if opaque else TF_HAS_ALPHA)
    pts = []
    cams = np.asarray(cams, dtype=np.float32).reshape(-1, 6)
    h, w = z.shape[1:3]
    for i, c in enumerate(cams):
        depth = 32768.0 / max(int(z[i, h // 2, w // 2]), 1)  # the z buffer holds 0x8000 / depth
        p, y = np.radians(c[3]), np.radians(c[4])
        fwd = np.array([np.cos(p) * np.cos(y), np.cos(p) * np.sin(y), -np.sin(p)])
        pt = c[:3] + fwd * depth
        r.decal_shoot(tex, pt, 0, scale)
        pts.append(pt)
    return pts
