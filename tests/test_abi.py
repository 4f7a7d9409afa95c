"""The drop-in boundary: include/elb200.h is the contract, libelb200.so must export every
function it declares with C linkage, and the ctypes table in elemental_b200/_lib.py must cover
the same set with the same arity.  CPU-only: symbols are resolved, no compute entry is called."""
import ctypes as C
import os
import re
import subprocess

import pytest

from elemental_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "elb200.h")


def _declarations():
    """{name: number of parameters} for every function prototype in the header."""
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    text = re.sub(r"^\s*#.*$", " ", text, flags=re.M)
    out = {}
    for m in re.finditer(r"\b(elb200_[a-z0-9_]+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        name, args = m.group(1), " ".join(m.group(2).split())
        out[name] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def test_header_declares_the_hot_path():
    d = _declarations()
    for must in ("elb200_gemm_f32", "elb200_gemm_f64", "elb200_copy2d_f32", "elb200_transpose_f64",
                 "elb200_axpy2d_f32", "elb200_scale2d_f64", "elb200_fill2d_f32", "elb200_lattice_f32",
                 "elb200_plan_copy", "elb200_plan_contract", "elb200_grid_create", "elb200_dm_create",
                 "elb200_dm_copy", "elb200_dm_axpy_contract", "elb200_dm_local_gemm", "elb200_dist_gemm"):
        assert must in d, must
    assert len(d) >= 50


def test_every_declared_symbol_is_exported():
    missing = [n for n in _declarations() if not hasattr(L.lib, n)]
    assert not missing, f"declared in include/elb200.h but not exported by libelb200.so: {missing}"


def test_exports_have_c_linkage_and_nothing_undeclared():
    """`nm -D` view: the elb200_* dynamic symbols are exactly the header's set (unmangled)."""
    out = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("elb200_")}
    declared = set(_declarations())
    assert declared <= exported
    assert exported <= declared, f"exported but not declared in the header: {sorted(exported - declared)}"


def test_ctypes_table_matches_header():
    d = _declarations()
    assert set(L.SIGNATURES) == set(d), (sorted(set(d) - set(L.SIGNATURES)), sorted(set(L.SIGNATURES) - set(d)))
    for name, (_res, args) in L.SIGNATURES.items():
        assert len(args) == d[name], f"{name}: header has {d[name]} parameters, ctypes table {len(args)}"


def test_struct_layouts_match_header_sizes():
    """Plain-C structs crossing the boundary: sizes implied by the header's field lists."""
    assert C.sizeof(L.Segment) == 7 * 8
    assert C.sizeof(L.DistSpec) == 2 * 4 + 4 * 8 + 8  # + int root, padded to the 8-byte alignment
    assert C.sizeof(L.Message) == 8 + C.sizeof(L.Segment)
    assert L.MAX_SEGMENTS == int(re.search(r"#define ELB200_MAX_SEGMENTS (\d+)", open(HEADER).read()).group(1))


def test_no_torch_types_cross_the_boundary():
    text = open(HEADER).read()
    assert 'extern "C"' in text
    for banned in ("at::", "torch::", "c10::", "std::", "#include <torch"):
        assert banned not in text, banned


def test_status_strings_and_build_info():
    assert L.lib.elb200_status_string(0).decode().lower().startswith("ok") or L.lib.elb200_status_string(0)
    info = L.lib.elb200_build_info().decode()
    assert "sm_100a" in info and "tcgen05" in info


def test_library_is_built_for_sm_100a_only():
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", L.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_hot_kernels_use_the_blackwell_paths():
    """Static check of the shipped binary: the fp32 GEMM issues tcgen05.mma (SASS UTCHMMA, both CTA-group
    forms) fed by TMA (UTMALDG) with mbarriers (SYNCS), the fp64 GEMM runs on the DMMA pipe, and the
    BLAS-1 / pack kernels use 128-bit global accesses — and nothing fell back to mma.sync-era HMMA or to
    a library GEMM (libelb200.so must not link cuBLAS)."""
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", L.LIB_PATH], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA.2CTA", "UTCHMMA ", "UTMALDG.2D", "SYNCS.PHASECHK", "DMMA.8", "LDG.E.128", "STG.E.128"):
        assert mnemonic in sass, mnemonic
    assert not re.search(r"(?<![A-Z])HMMA\.", sass)  # mma.sync / wmma half-precision path (UTCHMMA is tcgen05)
    needed = subprocess.run(["readelf", "-d", L.LIB_PATH], capture_output=True, text=True).stdout
    assert "cublas" not in needed.lower(), needed


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: nothing under elemental_b200/ may import or execute it."""
    pkg = os.path.join(ROOT, "elemental_b200")
    for dirpath, _dirs, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".hpp", ".h", ".cuh")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(dirpath, f)
                assert "oracle/" not in src, os.path.join(dirpath, f)
