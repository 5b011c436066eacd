"""The drop-in boundary (not gpu): libref_cuda.so loads, exports every symbol include/ref_cuda.h declares and
nothing else, completes the GetRefAPI handshake, and R_Init fails loudly when no CUDA device is present
(there is no CPU fallback behind the plugin)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ref_cuda.h")).read()
    src = src.replace("#define REF_CUDA_EXPORT", "")
    return sorted(set(re.findall(r"^REF_CUDA_EXPORT[^;(]*?\b(\w+)\s*\(", src, re.M)))


def test_header_declares_the_reference_entry_point():
    syms = declared_symbols()
    assert "GetRefAPI" in syms and "RefCuda_RenderViews" in syms and len(syms) >= 6


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.REF_CUDA_PATH)
    for s in declared_symbols():
        assert hasattr(lib, s), s


def test_library_exports_nothing_else(pkg):
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.REF_CUDA_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert exported == declared_symbols()


def test_product_library_does_not_link_the_oracle(pkg):
    out = subprocess.run(["ldd", pkg.REF_CUDA_PATH], capture_output=True, text=True).stdout
    assert "ref_soft" not in out and "emu" not in out
    syms = subprocess.run(["nm", pkg.REF_CUDA_PATH], capture_output=True, text=True).stdout
    assert "rc_walk_world" in syms or "k_viswalk" in syms  # the kernels are in there
    assert "R_ScanEdges" not in syms  # none of ref_soft is


def test_version_mismatch_is_rejected(pkg):
    """ref/common/ref_context.c:60-61: GetRefAPI returns 0 for a foreign API version."""
    lib = ctypes.CDLL(pkg.REF_CUDA_PATH)
    lib.GetRefAPI.restype = ctypes.c_int
    buf = ctypes.create_string_buffer(8192)
    assert lib.GetRefAPI(17, buf, buf, buf) == 0


def test_init_fails_loudly_without_a_gpu(pkg, scene):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(RuntimeError, match="R_Init failed"):
        scene.Renderer(pkg.REF_CUDA_PATH, 320, 240)
