"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports everything include/cnrf.h declares,
the ctypes table covers it, and the product path refuses to run without CUDA (no CPU fallback)."""
import os
import re

import numpy as np
import pytest
import torch

from codenerf_b200 import _lib, ops

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "cnrf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cnrf_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"libcnrf.so does not export {n}"
    assert sorted(_lib.EXPORTS) == names, "ctypes signature table and include/cnrf.h disagree"
    assert lib.cnrf_abi_version() == 1
    assert lib.cnrf_error_string(-3).decode().startswith("workspace")


def test_size_queries_run_without_gpu():
    lib = _lib.load()
    import ctypes as C
    shp = _lib.FieldShape(256, 256, 63, 27, 4096, 160, 0, _lib.PRECISION_FP32, 1)
    saved = lib.cnrf_codenerf_saved_bytes(C.byref(shp))
    M = 4096 * 160
    assert saved >= 4 * (4 * M * 256 + M * 257)
    assert lib.cnrf_field_workspace_bytes(C.byref(shp), 1) > lib.cnrf_field_workspace_bytes(C.byref(shp), 0) > 0
    bad = _lib.FieldShape(256, 256, 63, 27, 0, 160, 0, 0, 1)
    assert lib.cnrf_codenerf_saved_bytes(C.byref(bad)) == 0


def test_no_cpu_fallback():
    x = torch.zeros(4, 3)
    with pytest.raises(_lib.CnrfError):
        ops.pos_embed(x, torch.ones(2), True)
    with pytest.raises(_lib.CnrfError):
        ops.composite(torch.zeros(2, 4, 4), torch.zeros(2, 4), torch.ones(2, 3))
    import view_synthesis.models as models
    m = models.CodeNeRFModel(64, 1, 64, 64, 10, 4, True, True)
    with pytest.raises(_lib.CnrfError):
        m(torch.zeros(3, 64), torch.zeros(3, 64), torch.zeros(3, 90))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "code-nerf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f"{f} imports the oracle"
                assert "codenerf_oracle" not in txt, f"{f} references the oracle"
