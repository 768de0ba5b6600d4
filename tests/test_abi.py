"""The C-ABI library builds, loads WITHOUT a GPU and exports every symbol the header declares; the
ctypes table names exactly the same set.  No compute calls here."""
import ctypes as C
import os
import re

from butterfly_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "butterfly_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(bfly_[a-z0-9_]+)\s*\(", text))


def test_header_and_ctypes_table_agree():
    hdr = header_symbols()
    assert len(hdr) >= 25
    assert hdr == set(_lib.SIGNATURES), (hdr ^ set(_lib.SIGNATURES))


def test_integration_notes_name_every_entry_point():
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = sorted(s for s in header_symbols() if s not in doc)
    assert not missing, f"INTEGRATION.md does not mention {missing}"


def test_library_loads_and_exports_every_symbol():
    lib = _lib.load()
    for name in header_symbols():
        assert hasattr(lib, name), name
    assert lib.bfly_version() >= 100


def test_errors_without_gpu_are_reported_not_crashes():
    import torch
    lib = _lib.load()
    if torch.cuda.is_available():
        return
    desc = _lib.ModelDesc(family=1, hidden=256, n_layers=1, n_heads=8, n_kv_heads=4, head_dim=32, ffn=512,
                          vocab=100, norm_eps=1e-5, rope_theta=1e4, max_pos=64, tp=1, rank=0, page_size=32,
                          max_batch=1, max_ctx=32, max_tokens=32)
    ctx = C.c_void_p()
    rc = lib.bfly_create(C.byref(desc), 0, C.byref(ctx))
    assert rc != 0 and not ctx.value
    assert b"no CUDA device" in lib.bfly_last_error(None)
    assert lib.bfly_decode(None, None, 1, None, None) != 0          # null ctx is an error, not a crash
    import pytest
    from butterfly_b200 import Model, SyntheticSource, make_plan, tiny
    with pytest.raises(RuntimeError):
        Model(make_plan(tiny("llama"), 1), 0, SyntheticSource(tiny("llama")))
