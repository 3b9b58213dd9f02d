"""The C-ABI library loads on a machine without a GPU, exports every symbol include/scenesubmit.h declares,
and refuses to create a context when no CUDA device exists (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from vulkanengine_b200 import scenesubmit as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "scenesubmit.h")).read()
    return sorted(set(re.findall(r"SS_API\s+[\w\s\*]+?\b(ss_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(S.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = S.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), f"libscenesubmit.so does not export {name}"
    assert lib.ss_version() == (0 << 16) | 1


def test_struct_sizes_match_header():
    assert C.sizeof(S.Config) == 48
    assert S.SUBMESH_DTYPE.itemsize == 112 and S.VIEW_DTYPE.itemsize == 68 and S.DRAW_DTYPE.itemsize == 20


def test_product_never_imports_the_oracle():
    """Only tests/, smoke() and bench.py's cpu_baseline legs may touch oracle/."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "vulkanengine_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "scenesubmit_oracle" not in text and "liboracle" not in text, f


def test_create_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(S.SceneSubmitError) as e:
        S.SceneSubmit(16, 16, 4)
    assert e.value.status == S.SS_ERR_CUDA


def test_sort_plan_scratch_is_monotone():
    lib = S.load_library()
    sizes = [lib.ss_sort_scratch_bytes(C.c_uint64(n)) for n in (0, 1, 4096, 4097, 1 << 20, 1 << 24)]
    assert sizes == sorted(sizes) and sizes[0] > 0
