"""The C-ABI library loads on a machine with no GPU, exports every symbol include/ptb200.h
declares, and fails loudly (no CPU fallback) when asked to compute without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from ptb200 import capi, device, scenes
from ptb200.builder import product_builder


def test_header_symbols_exported():
    header = open(os.path.join(capi.REPO_ROOT, "include", "ptb200.h")).read()
    declared = sorted(set(re.findall(r"\b(ptb200_[a-z0-9_]+)\s*\(", header)))
    assert declared == sorted(capi.ABI_SYMBOLS)
    lib = capi.product()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ptb200_abi_version() == 2


def test_struct_layouts_match_header():
    assert C.sizeof(capi.PtbPrim) == 32 and C.sizeof(capi.PtbXform) == 96 and C.sizeof(capi.PtbMaterial) == 32
    assert C.sizeof(capi.PtbRenderParams) == 72 and C.sizeof(capi.PtbStats) == 96


def test_product_does_not_link_the_oracle():
    """The shipped library must not depend on anything under oracle/."""
    out = os.popen("ldd %s" % capi.PRODUCT_LIB).read()
    assert "ptoracle" not in out and "ptref" not in out
    syms = os.popen("nm -D %s" % capi.PRODUCT_LIB).read()
    assert " orc_" not in syms and " ref_" not in syms


def test_no_device_fails_loudly():
    lib = capi.product()
    if lib.ptb200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    flat = device.Flat(scenes.config2(product_builder()))
    with pytest.raises(capi.PtbError) as e:
        device.DeviceScene(flat)
    assert e.value.status == -4 and "no CPU fallback" in str(e.value)


def test_malformed_programs_are_rejected():
    lib = capi.product()
    flat = device.Flat(scenes.config2(product_builder()))
    d = capi.PtbSceneDesc.from_buffer_copy(flat.desc)
    handle = C.c_void_p()
    d.abi_version = 99
    assert lib.ptb200_scene_create(C.byref(d), C.byref(handle)) == -1
    assert b"abi_version" in lib.ptb200_last_error()
    if lib.ptb200_device_count() == 0:
        return
    d = capi.PtbSceneDesc.from_buffer_copy(flat.desc)
    d.n_ops -= 1  # leaves two objects on the stack
    assert lib.ptb200_scene_create(C.byref(d), C.byref(handle)) == -1
