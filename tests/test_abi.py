"""The C-ABI library loads and exports every symbol include/crt.h declares (no compute, no GPU)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from computeraytracer_b200.build import build_libcrt
    return C.CDLL(build_libcrt())


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "crt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"CRT_API\s+[\w\s\*]+?\b(crt_\w+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_symbols()
    for must in ("crt_create", "crt_destroy", "crt_set_binding", "crt_reset", "crt_set_params", "crt_dispatch_frames",
                 "crt_read_binding", "crt_trace_rays", "crt_get_counters", "crt_set_shard", "crt_device_ptr",
                 "crt_last_error", "crt_version"):
        assert must in names


def test_every_declared_symbol_is_exported(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_version_string(lib):
    lib.crt_version.restype = C.c_char_p
    assert b"sm_100a" in lib.crt_version()


def test_default_params_are_the_reference_literals(lib):
    from computeraytracer_b200.renderer import Params
    p = Params()
    lib.crt_default_params.argtypes = [C.POINTER(Params)]
    lib.crt_default_params.restype = None
    lib.crt_default_params(C.byref(p))
    # ComputeShader.wgsl:14, :15, :720, :112
    assert (p.max_depth, p.grid_size) == (100, 16)
    assert abs(p.t_min - 0.001) < 1e-9 and abs(p.exposure - 2.2) < 1e-6


def test_no_cpu_fallback(lib):
    """Without a B200 the product path fails loudly instead of computing on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    lib.crt_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    assert lib.crt_create(0, C.byref(h)) == 7  # CRT_ERR_NODEVICE
    assert not h.value
    from computeraytracer_b200.renderer import CrtError, Renderer
    with pytest.raises(CrtError):
        Renderer(0)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may reference it."""
    pkg = os.path.join(ROOT, "computeraytracer_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dp, f), errors="ignore").read()
                assert "crt_oracle" not in text and "libcrt_oracle" not in text and "orc_" not in text, os.path.join(dp, f)
