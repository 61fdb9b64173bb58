"""The C-ABI library loads on a CPU-only box and exports every symbol include/mggcn.h declares."""
import ctypes
import os
import re


def declared_symbols(header):
    txt = open(header).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"MGGCN_API\s+[\w\s\*]+?\b(mggcn_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(mg):
    names = declared_symbols(mg.INCLUDE)
    assert len(names) > 40
    L = ctypes.CDLL(mg.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # and the ctypes table used by tests/bench covers the same set
    assert sorted(mg.exported_symbols()) == names


def test_abi_version_and_error_text(mg):
    assert mg.lib().mggcn_abi_version() == 1
    rc = mg.lib().mggcn_tune(b"no-such-knob", 1)
    assert rc == -1
    assert b"no-such-knob" in mg.lib().mggcn_last_error()


def test_no_oracle_in_product(mg):
    """The product tree must not reference oracle/ (SURVEY 8c / tier rule 3)."""
    root = os.path.dirname(mg.LIB_PATH)
    pkg = os.path.dirname(root)
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath or dirpath.endswith("lib"):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in txt and "orc_" not in txt, f


def test_device_entry_points_fail_loudly_without_gpu(mg):
    import torch
    if torch.cuda.is_available():
        return
    h = ctypes.c_void_p()
    import numpy as np
    rp = np.array([0, 1, 2], np.uint32)
    rc = mg.lib().mggcn_spmm_plan_create_i32(2, rp.ctypes.data, 0, ctypes.byref(h))
    assert rc != 0  # a CUDA error code, not a silent CPU path
