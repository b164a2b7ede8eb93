"""The C-ABI library loads on a CPU-only box and exports every symbol include/k8sgpu.h declares.
No compute calls here (no GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def capi():
    import __graft_entry__ as g

    g.build()
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi as c

    return c


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "k8sgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(k8sgpu_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree(capi):
    assert declared_symbols() == sorted(capi.EXPORTS)


def test_library_exports_every_declared_symbol(capi):
    lib = ctypes.CDLL(capi.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), name
    assert lib.k8sgpu_abi_version() == 2


def test_struct_layouts_match_header(capi):
    # sizes a cgo shim would see on LP64
    assert ctypes.sizeof(capi.NodeTable) == 8 + 8 * 8
    assert ctypes.sizeof(capi.PodTable) == 8 + 8 * 8
    assert ctypes.sizeof(capi.PoolTable) == 24
    assert ctypes.sizeof(capi.VmTable) == 32
    assert ctypes.sizeof(capi.ReconcileOut) == 12 * 8
    assert ctypes.sizeof(capi.Policy) == 8
    assert ctypes.sizeof(capi.PlaceOut) == 32
    assert ctypes.sizeof(capi.Stats) == 64


def test_no_gpu_means_error_not_fallback(capi):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.K8sGpuError) as e:
        capi.Context(0)
    assert e.value.code == capi.E_CUDA


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "k8s-gpu-_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp", ".go")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in src.lower(), os.path.join(dirpath, f)
