"""The C-ABI library loads on a machine without a GPU and exports every symbol include/efl.h declares."""
import ctypes
import os
import re

from tests.conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "efl.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(efl_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib_path = os.path.join(ROOT, "singa_b200", "libefl.so")
    assert os.path.isfile(lib_path), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(lib_path)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/efl.h but not exported"


def test_python_binding_covers_header():
    from singa_b200 import _lib
    bound = set(_lib.SIGNATURES) | {"efl_version", "efl_last_error_string", "efl_gemm_tn_splits", "efl_device_cc",
                                    "efl_tc_gemm_tn_splits", "efl_tc_gemm_pair_mode", "efl_tc_gemm_ring_kb", "efl_rotate_back_parts", "efl_radial_mlp_part_size", "efl_radius_graph_temp_bytes", "efl_csr_temp_bytes", "efl_radial_hidden_part_size"}
    assert set(declared_symbols()) <= bound, set(declared_symbols()) - bound
    assert _lib.LIB.efl_version() >= 100


def test_ops_refuse_cpu_tensors():
    import pytest
    import torch
    from singa_b200 import _lib, ops
    with pytest.raises(_lib.EflError):
        ops.wigner_from_rotmat(torch.eye(3).repeat(4, 1, 1), 6, 2)


def test_python_signatures_match_the_header_prototypes():
    """ctypes argument codes of singa_b200/_lib.py against the C prototypes of include/efl.h, argument by argument
    (pointer / int / long long / float): a drifted binding corrupts the call stack instead of failing loudly."""
    from singa_b200 import _lib
    text = open(os.path.join(ROOT, "include", "efl.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    checked = 0
    for m in re.finditer(r"int\s+(efl_\w+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        name, args = m.group(1), m.group(2)
        if name not in _lib.SIGNATURES:
            continue
        codes = ""
        for a in (x.strip() for x in args.split(",")):
            codes += "p" if "*" in a else ("l" if a.startswith("long long") else ("f" if a.startswith("float") else "i"))
        assert codes == _lib.SIGNATURES[name], (name, codes, _lib.SIGNATURES[name])
        checked += 1
    assert checked >= 35
