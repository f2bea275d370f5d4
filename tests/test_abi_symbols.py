"""The C-ABI libraries load without a GPU and export every symbol the public headers declare
(include/b200_kernels.h -> libb200_kernels.so, include/b200_backend.h -> libggml-b200.so).  No compute calls."""
import ctypes as C
import os
import re

import pytest

from conftest import REF_DIR, ROOT

PKG = os.path.join(ROOT, "llama-mi50.cpp_b200")


def _declared(header, pattern):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(pattern, src)))


def test_kernel_library_exports_every_declared_symbol():
    names = _declared("b200_kernels.h", r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(")
    assert len(names) >= 30, names
    lib = C.CDLL(os.path.join(PKG, "libb200_kernels.so"))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/b200_kernels.h but not exported: {missing}"
    # validation mode works without a device: an obviously bad call is rejected, not executed
    lib.b200_gemv_fused_supported.argtypes = [C.c_void_p]
    assert lib.b200_gemv_fused_supported(None) == 0
    lib.b200_mmq_supported.argtypes = [C.c_void_p]
    assert lib.b200_mmq_supported(None) == 0


def test_plugin_exports_the_loader_entry_points():
    so = os.path.join(PKG, "libggml-b200.so")
    base = os.path.join(REF_DIR, "libggml-base.so")
    if not (os.path.exists(so) and os.path.exists(base)):
        pytest.skip("plugin / oracle/_ref not built (needs /root/reference at build time)")
    C.CDLL(base, mode=C.RTLD_GLOBAL)          # the host process (llama-bench, test-backend-ops) provides ggml-base
    lib = C.CDLL(so)
    names = _declared("b200_backend.h", r"\b(ggml_backend_\w+)\s*\(")
    assert {"ggml_backend_init", "ggml_backend_score"} <= set(names)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/b200_backend.h but not exported: {missing}"
    lib.ggml_backend_score.restype = C.c_int
    assert lib.ggml_backend_score() == 0      # no sm_100 device in this container: the loader must skip the plugin
