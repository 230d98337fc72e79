"""The C-ABI library loads on a CPU-only box and exports every symbol that
include/qeb200.h declares; compute entry points fail loudly without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g

    if not os.path.exists(os.path.join(ROOT, "q-e_b200", "libqeb200.so")):
        g.build()
    import qe_b200

    return qe_b200


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "qeb200.h")).read()
    names = sorted(set(re.findall(r"\b(qeb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 20
    L = ctypes.CDLL(lib.lib_path())
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/qeb200.h but not exported"


def test_divide_all_matches_reference_rule(lib):
    from oracle import hpsi_ref as R

    for m in (0, 1, 7, 10, 800):
        for nproc in (1, 2, 3, 4, 8):
            counts, displs = R.divide_all(m, nproc)
            for me in range(nproc):
                s, e = lib.divide_all(m, nproc, me)
                assert (s, e - s) == (int(displs[me]), int(counts[me]))


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.QebError):
        lib.HPsi()


def test_fortran_shim_interfaces_match_header():
    """fortran/qeb200_shim.f90 cannot be compiled here (no Fortran compiler): check at least that every BIND(C) interface
    it declares names a function of include/qeb200.h with the same number of arguments, and that every C call in the
    shim's executable part has an interface."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "qeb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    cdecl = {}
    for mm in re.finditer(r"\b(qeb_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", hdr, flags=re.S):
        args = mm.group(2).strip()
        cdecl[mm.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    shim = open(os.path.join(root, "fortran", "qeb200_shim.f90")).read()
    shim = re.sub(r"&\s*\n\s*&?", " ", shim)  # join continuation lines
    found = {}
    for mm in re.finditer(r"FUNCTION\s+(\w+)\s*\(([^)]*)\)\s*BIND\(C,\s*name='(\w+)'\)", shim, flags=re.I):
        fname, args, cname = mm.group(1), mm.group(2), mm.group(3)
        assert fname == cname
        found[cname] = len([a for a in args.split(",") if a.strip()])
    assert len(found) > 20
    for name, nargs in found.items():
        assert name in cdecl, f"{name}: interface in the shim, not in the header"
        assert cdecl[name] == nargs, f"{name}: {nargs} arguments in the shim, {cdecl[name]} in the header"
    body = re.sub(r"!.*", "", shim)
    called = set(re.findall(r"\b(qeb_[a-z0-9_]+)\s*\(", body))
    for name in called:
        assert name in found, f"{name} is called in the shim without an interface"
