"""CPU-side checks of the C-ABI boundary: the library builds, loads without a GPU, and exports every symbol that
include/ufm_b200.h declares.  No compute calls here."""
import os
import re

import pytest

import __graft_entry__ as g
from ufm_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    g.build()
    return _lib.load()


def test_header_symbols_exported(lib, repo_root):
    hdr = open(os.path.join(repo_root, "include", "ufm_b200.h")).read()
    declared = set(re.findall(r"\b(ufm_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"ufm_engine", "ufm_config"}
    assert len(declared) >= 18
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/ufm_b200.h but not exported"
    assert declared == set(_lib.SYMBOLS), (declared ^ set(_lib.SYMBOLS))


def test_create_fails_loudly_without_gpu(lib):
    import ctypes as C

    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    cfg = _lib.UfmConfig()
    h = C.c_void_p()
    assert lib.ufm_create(C.byref(cfg), 0, C.byref(h)) != 0
    assert not h.value


def test_engine_refuses_cpu():
    import torch

    from ufm_b200 import UniFlowMatchConfidence, get_config

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    m = UniFlowMatchConfidence(**get_config("tiny_base"))
    src = torch.zeros(1, 168, 224, 3, dtype=torch.uint8)
    with pytest.raises(RuntimeError, match="CUDA only"):
        m.predict_correspondences_batched(src, src)
