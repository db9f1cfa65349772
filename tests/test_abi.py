"""The C-ABI library loads and exports every symbol include/dls_b200.h declares (no compute)."""
import os
import re

import dls_b200
from dls_b200 import _lib

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(REPO, "include", "dls_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dls_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    syms = declared_symbols()
    assert "dls_agg_spmm_f32" in syms and "dls_train_wave" in syms and "dls_eval_wave" in syms
    l = _lib.lib()
    for s in syms:
        assert hasattr(l, s), "library does not export %s" % s
        assert s in _lib.SIGNATURES, "no ctypes signature for %s" % s
    assert sorted(_lib.SIGNATURES) == syms


def test_version_and_sizes():
    l = _lib.lib()
    assert l.dls_version() >= 100
    assert b"sm_100a" in l.dls_build_info()
    n = l.dls_train_workspace_bytes(0, 256, 32)
    assert n > 0 and n % 256 == 0
    assert l.dls_train_workspace_bytes(0, 4, 65) == -3       # batch above DLS_MAX_GROUP
    assert l.dls_eval_workspace_bytes(0, 16, 2048) > 0


def test_argument_errors_without_gpu():
    l = _lib.lib()
    assert l.dls_agg_spmm_f32(None, 0, None, 0, None, None, None, None, None, 1, 1, 1, 4, 0, None) == -1
    assert l.dls_copy_rows(None, 0, None, 0, None, None, 1, 4, None) == -1
