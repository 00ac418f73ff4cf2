"""CPU-side checks of the drop-in boundary: liblora_phy_b200.so loads without a GPU, exports exactly the
entry points include/lora_phy_b200.h declares, and fails loudly (no CPU fallback) when asked to compute
without a device."""
import ctypes as C
import os
import re

import pytest

import lora_phy_b200 as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "lora_phy_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lora_b200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = L.load_library()
    names = declared_functions()
    assert len(names) == 38
    for n in names:
        assert getattr(lib, n) is not None, n
    assert sorted(L.EXPORTS) == names  # the Python layer binds all of them, and nothing undeclared


def test_struct_layouts_match_header():
    assert L.PKT_METRICS.itemsize == 16 and L.CHAIN_RESULT.itemsize == 16
    assert C.sizeof(L.Config) == 20 and C.sizeof(L.ChainCfg) == 40


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the failure path is for CPU-only hosts")
    with pytest.raises(L.LoraB200Error) as e:
        L.Context(7)
    assert "-2" in str(e.value)  # LORA_B200_ECUDA


def test_argument_guards_need_no_device():
    lib = L.load_library()
    h = C.c_void_p()
    assert lib.lora_b200_ctx_create(None, C.byref(h)) == -1
    bad = L.Config(13, 125000, 1, 0, -1)
    assert lib.lora_b200_ctx_create(C.byref(bad), C.byref(h)) == -1
    bad = L.Config(7, 100000, 1, 0, -1)
    assert lib.lora_b200_ctx_create(C.byref(bad), C.byref(h)) == -1
    assert b"bw_hz" in lib.lora_b200_last_error()
    assert lib.lora_b200_demod_batch(None, 0, None, 0, 0, 0, None, 0, None, None, None) == -1
