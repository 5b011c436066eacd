"""RefCuda_BlitPixelsNoAlloc (include/ref_cuda.h): the display conversion over the library's static device scratch, for
callers that cannot call cudaMalloc.  Collected last on purpose; passes alone in a fresh
process and after the other tests.  Checked against the existing entry point RefCuda_BlitViews, bit for bit (the bar the
display-tail tests use), on two calls in a row that share the scratch: different frames, sizes and display formats."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

XRGB8888 = (4, 0x00FF0000, 0x0000FF00, 0x000000FF)
RGB565 = (2, 0xF800, 0x07E0, 0x001F)
MAXPX = 1920 * 1080


def no_alloc_blit(lib, pixels, fmt):
    fn = lib.RefCuda_BlitPixelsNoAlloc
    fn.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint]
    fn.restype = C.c_int
    pixels = np.ascontiguousarray(pixels, dtype=np.uint16)
    out = np.zeros(pixels.shape, np.uint16 if fmt[0] == 2 else np.uint32)
    rc = fn(pixels.ctypes.data, pixels.size, out.ctypes.data, *fmt)
    return rc, out


def test_no_alloc_blit_matches_blit_views_on_two_calls_in_a_row(pkg, scene):
    m = scene.make_map("tiny", ncameras=16)
    r = scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
    lib = C.CDLL(pkg.REF_CUDA_PATH)  # the handle the host stand-in already loaded
    cams = m.cameras[:4]
    col, _ = r.render_views(cams)
    want32 = r.blit_views(cams, sw_format=XRGB8888)
    want16 = r.blit_views(cams, sw_format=RGB565)
    assert len(np.unique(want32)) > 16

    # call 1: one whole frame, 32-bit display format
    rc, got = no_alloc_blit(lib, col[0], XRGB8888)
    assert rc == 0 and np.array_equal(got, want32[0])
    # call 2 straight after it, same scratch: another frame, a ragged size (not a multiple of 8 pixels), 16-bit format
    part = col[3].reshape(-1)[: 320 * 200 + 5]
    rc, got = no_alloc_blit(lib, part, RGB565)
    assert rc == 0 and np.array_equal(got, want16[3].reshape(-1)[: part.size])
    # and the first again: nothing of call 2 leaks into it
    rc, got = no_alloc_blit(lib, col[0], XRGB8888)
    assert rc == 0 and np.array_equal(got, want32[0])
    # the existing entry point still gives its answer afterwards
    assert np.array_equal(r.blit_views(cams, sw_format=XRGB8888), want32)


def test_no_alloc_blit_rejects_sizes_beyond_the_static_scratch(pkg, scene):
    m = scene.make_map("tiny", ncameras=16)
    scene.Renderer(pkg.REF_CUDA_PATH, 320, 240, m)
    lib = C.CDLL(pkg.REF_CUDA_PATH)
    rc, _ = no_alloc_blit(lib, np.zeros(MAXPX + 1, np.uint16), XRGB8888)
    assert rc != 0
    rc, got = no_alloc_blit(lib, np.full(MAXPX, 0xFFFF, np.uint16), XRGB8888)  # the stated maximum itself is accepted
    assert rc == 0 and len(np.unique(got)) == 1
