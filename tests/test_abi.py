"""CPU-only: the C-ABI library loads and exports every symbol that
include/minnow_cuda.h declares; struct sizes match the reference layouts."""
import ctypes as C
import os
import re

import numpy as np

import minnow_c_b200 as mc


def declared_functions():
    text = open(os.path.join(mc.ROOT, "include", "minnow_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b((?:mnwcuda_|MnwCuda)\w+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    mc.build()
    handle = C.CDLL(mc.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 18
    for name in names:
        assert hasattr(handle, name), name
    assert set(names) == set(mc.EXPORTS), set(names) ^ set(mc.EXPORTS)


def test_size_functions_need_no_gpu():
    lib = mc.lib()
    assert lib.mnwcuda_FieldFixedBytes(0) == 64
    assert lib.mnwcuda_FieldFixedBytes(1) == 64 + 48
    assert lib.mnwcuda_FieldFixedBytes(1025) == 64 + 48 * 2
    assert lib.mnwcuda_FieldBound(1024) == 64 + 48 + 3 * 3200
    import _oracle
    o = _oracle.port()
    for n in (0, 1, 1024, 1025, 4194304):
        assert lib.mnwcuda_FieldBound(n) == o.lib.orc_diff_bound(n)
        assert lib.mnwcuda_FieldFixedBytes(n) == o.lib.orc_diff_fixed_bytes(n)


def test_struct_layouts():
    # sizes on x86-64 of /root/reference/src/types.h structs
    assert C.sizeof(mc.FieldHeader) == 16
    assert C.sizeof(mc.Field) == 40 and C.sizeof(mc.QField) == 40
    assert C.sizeof(mc.CField) == 40
    assert C.sizeof(mc.PositionAccuracy) == 24 and C.sizeof(mc.VelocityAccuracy) == 24
    assert C.sizeof(mc.PositionQuantization) == 48 and C.sizeof(mc.VelocityQuantization) == 48
    assert C.sizeof(mc.FieldDesc) == 32
    import _oracle
    assert _oracle.HEADER_DTYPE.itemsize == 64 and _oracle.ENTRY_DTYPE.itemsize == 16


def test_nccl_driver_exports_every_declared_symbol():
    """libminnow_nccl.so (include/minnow_nccl.h): symbols only -- loading it needs libnccl, which this image has"""
    mc.build()
    path = mc.build_nccl()
    text = open(os.path.join(mc.ROOT, "include", "minnow_nccl.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = sorted(set(re.findall(r"\b(mnwnccl_\w+)\s*\(", text)))
    assert set(names) == set(mc.NCCL_EXPORTS)
    import subprocess
    syms = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True).stdout
    for name in names:
        assert (" T " + name) in syms, name
