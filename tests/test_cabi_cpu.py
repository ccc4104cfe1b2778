"""CPU: the C-ABI library loads and exports every symbol include/dqn_b200.h declares (no compute calls)."""
import os
import re

from conftest import ROOT


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    import _cabi
    hdr = open(os.path.join(ROOT, "include", "dqn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(dqn_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(_cabi.lib, name), "libdqn_b200.so does not export %s" % name
        assert name in _cabi.SIGNATURES, "no ctypes signature for %s" % name
    assert set(_cabi.SIGNATURES) == declared
    assert _cabi.lib.dqn_version() >= 100


def test_ctypes_signatures_have_the_arity_of_the_header():
    import __graft_entry__ as ge
    ge.build()
    import _cabi
    hdr = open(os.path.join(ROOT, "include", "dqn_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    decls = re.findall(r"\b(dqn_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", hdr)
    assert len(decls) >= 25
    for name, params in decls:
        params = params.strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(_cabi.SIGNATURES[name][1]), "%s: header has %d parameters, _cabi.py %d" % (name, n, len(_cabi.SIGNATURES[name][1]))


def test_no_cuda_device_fails_loudly_not_silently():
    import torch
    if torch.cuda.is_available():
        return
    import __graft_entry__ as ge
    ge.build()
    import _cabi as c
    h = c.c_vp()
    rc = c.lib.dqn_replay_create(c.C.byref(h), 8, 0, 16, 1, 1, 0)
    assert rc != 0 and c.last_error()
