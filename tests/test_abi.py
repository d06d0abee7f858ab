"""The C-ABI shared library loads and exports exactly what include/cnx_b200.h declares.
No compute calls here (CPU box)."""
import os
import re

from conftest import ROOT


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "cnx_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cnx_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_bound_symbols():
    from image_classification_b200 import _lib
    assert _header_symbols() == sorted(_lib.SIGNATURES)


def test_library_exports_every_header_symbol(cnx_lib):
    for name in _header_symbols():
        assert hasattr(cnx_lib, name), name


def test_version_and_chunk(cnx_lib):
    assert cnx_lib.cnx_version() >= 100
    assert cnx_lib.cnx_ema_chunk_elems() == 8192


def test_status_enum_matches_header():
    text = open(os.path.join(ROOT, "include", "cnx_b200.h")).read()
    vals = dict(re.findall(r"(CNX_(?:OK|ERR_[A-Z0-9_]+))\s*=\s*(-?\d+)", text))
    assert vals["CNX_OK"] == "0" and len(vals) == 7
    assert len(set(vals.values())) == len(vals)


def test_plan_entry_layout_matches_header():
    import ctypes
    from image_classification_b200 import _lib
    from image_classification_b200.mix import PLAN_DTYPE
    assert ctypes.sizeof(_lib.MixPlanEntry) == 32 == PLAN_DTYPE.itemsize
    assert [f[0] for f in _lib.MixPlanEntry._fields_] == list(PLAN_DTYPE.names)


def test_no_gpu_means_loud_failure(cnx_lib):
    """Without an sm_100 device every op must refuse (no CPU fallback)."""
    import torch
    if torch.cuda.is_available():
        return
    from image_classification_b200 import _lib
    assert cnx_lib.cnx_check_device() == -4
    assert "sm_100" in _lib.last_error()
