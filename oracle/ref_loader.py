"""Import the reference's own Python modules from /root/reference (THIS container only).

TEST INFRASTRUCTURE.  `timm` and `albumentations` are absent, so stub modules are inserted into
sys.modules before importing `dataset.py` / `train_advanced_v4.py` verbatim.  The
import is read-only with respect to /root/reference; it is done from a scratch cwd because the
training script opens a log file at import time (train_advanced_v4.py:25-32).

Nothing that runs on the GPU box may call this: /root/reference does not exist there.  It is used
by oracle/gen_golden.py (fixture generation) and by `-m "not gpu"` tests that skip when the
reference tree is missing.
"""
from __future__ import annotations

import importlib
import importlib.machinery
import os
import sys
import tempfile
import types

REFERENCE_ROOT = "/root/reference"


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "train_advanced_v4.py"))


def _stub(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, loader=None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


_cache: dict[str, types.ModuleType] = {}


def load_reference():
    """-> (dataset_module, train_module) of the reference, imported verbatim."""
    if "dataset" in _cache:
        return _cache["dataset"], _cache["train"]
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE_ROOT} is not present on this machine")
    if "timm" not in sys.modules:
        _stub("timm", create_model=None)
    if "albumentations" not in sys.modules:
        a = _stub("albumentations")
        ap = _stub("albumentations.pytorch", ToTensorV2=object)
        a.pytorch = ap
    cwd = os.getcwd()
    scratch = tempfile.mkdtemp(prefix="cnx_ref_")
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        os.chdir(scratch)
        ds = importlib.import_module("dataset")
        tr = importlib.import_module("train_advanced_v4")
    finally:
        os.chdir(cwd)
        sys.path.remove(REFERENCE_ROOT)
    _cache["dataset"], _cache["train"] = ds, tr
    return ds, tr
