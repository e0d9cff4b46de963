"""CPU suite: the C-ABI library loads, exports every symbol include/trajopt_cuda.h declares, and refuses to
compute without a GPU (no fallback); the product tree never reaches into oracle/."""
import ctypes as C
import os
import re
import subprocess

import pytest

from tests import helpers

HEADER = os.path.join(helpers.ROOT, "include", "trajopt_cuda.h")
PKG = os.path.join(helpers.ROOT, "trajopt-1_b200")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tc_[a-z0-9_]+)\s*\(", src)))


def lib():
    path = helpers.pkg.LIB_PATH
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(PKG, "csrc"), "-s", "-j4"])
    return C.CDLL(path)


def test_every_declared_symbol_is_exported():
    names = declared_functions()
    assert len(names) >= 29 and "tc_eval" in names and "tc_allgather_best" in names
    dll = lib()
    missing = [n for n in names if not hasattr(dll, n)]
    assert not missing, missing
    out = subprocess.run(["nm", "-D", "--defined-only", helpers.pkg.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r"\bT (tc_[a-z0-9_]+)$", out, flags=re.M))
    assert exported == set(names), (sorted(exported - set(names)), sorted(set(names) - exported))


def test_abi_version_and_record_sizes():
    dll = lib()
    assert dll.tc_abi_version() == 2
    capi = helpers.capi
    assert capi.CONTACT_DTYPE.itemsize == 200 and capi.GRADIENT_DTYPE.itemsize == 64 and capi.SET_DTYPE.itemsize == 112



def test_header_is_plain_c_and_matches_the_ctypes_mirror(tmp_path):
    """include/trajopt_cuda.h compiles as C99 on its own; struct sizes agree with trajopt-1_b200/capi.py"""
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "trajopt_cuda.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(tc_scene_desc),sizeof(tc_eval_desc),sizeof(tc_contact),sizeof(tc_gradient),sizeof(tc_counts),"
                   "sizeof(tc_set));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.dirname(HEADER), str(src),
                           "-o", str(exe)])
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    capi = helpers.capi
    assert sizes == [C.sizeof(capi.SceneDesc), C.sizeof(capi.EvalDesc), capi.CONTACT_DTYPE.itemsize,
                     capi.GRADIENT_DTYPE.itemsize, 8 * len(capi.COUNT_FIELDS), capi.SET_DTYPE.itemsize], sizes


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    dll = lib()
    dll.tc_last_error.restype = C.c_char_p
    assert dll.tc_init(0) == 2  # TC_ERR_CUDA
    assert len(dll.tc_last_error()) > 0
    # nothing computes before a successful tc_init
    s = helpers.load_scene("boxbot")
    with pytest.raises(helpers.capi.TcError):
        helpers.capi.Lib(helpers.pkg.LIB_PATH, "tc_").scene(s, 0.1, 1.0, 0.01)


def test_product_tree_does_not_touch_the_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may use oracle/"""
    hits = []
    for root, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(root, f), errors="ignore").read()
                if re.search(r"pyoracle|liboracle|oracle/|import oracle|from oracle", txt):
                    hits.append(os.path.join(root, f))
    assert not hits, hits


def test_noalloc_limit_matches_the_header_and_needs_no_device_to_reject():
    """TC_NOALLOC_MAX_TERMS of the header = the constant of the Python mirror; the size and argument checks of
    tc_cost_noalloc come before any CUDA call, so they answer without a GPU."""
    m = re.search(r"#define\s+TC_NOALLOC_MAX_TERMS\s+(\d+)", open(HEADER).read())
    assert m and int(m.group(1)) == helpers.capi.NOALLOC_MAX_TERMS
    dll = lib()
    dll.tc_last_error.restype = C.c_char_p
    one = C.c_void_p(8)  # never dereferenced: the call is rejected first
    args = lambda n, per: (one, C.c_int32(n), C.c_int32(per), one, one, one, one, C.c_void_p(0))
    assert dll.tc_cost_noalloc(*args(helpers.capi.NOALLOC_MAX_TERMS + 1, 1)) == helpers.capi.TC_ERR_OVERFLOW
    assert b"TC_NOALLOC_MAX_TERMS" in dll.tc_last_error()
    assert dll.tc_cost_noalloc(*args(7, 2)) == helpers.capi.TC_ERR_INVALID
    assert dll.tc_cost_noalloc(*args(0, 1)) == helpers.capi.TC_ERR_INVALID
