"""CPU: the C-ABI shared library loads and exports every symbol include/f5b200.h declares; without a GPU every compute
entry fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

from f5b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "f5b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(f5b_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/f5b200.h but not exported"
    assert sorted(_lib.EXPORTED) == names  # the ctypes table covers the whole header
    assert L.f5b_version() >= 100


@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-only behaviour")
def test_no_gpu_means_loud_failure():
    L = _lib.lib()
    assert L.f5b_device_check(0) == _lib.ERR_NO_DEVICE
    assert b"no CPU fallback" in L.f5b_last_error()
    h = C.c_void_p()
    arch = _lib.Arch(0, 128, 1, 2, 64, 2, 100, 10, 64, 0, 1, 0, 0)
    w = torch.zeros(4096)
    arr = (_lib.Tensor * 1)(_lib.Tensor(b"x", w.data_ptr(), 4, 0))
    assert L.f5b_plan_create(C.byref(arch), arr, 1, 0, 0, C.byref(h)) == _lib.ERR_NO_DEVICE
    with pytest.raises(_lib.F5BError):
        _lib.check(L.f5b_mel(w.data_ptr(), 1, 2048, w.data_ptr(), 0, 0, None))
    import f5b200

    with pytest.raises(_lib.F5BError):
        f5b200.MelSpec()(torch.zeros(1, 2048))
