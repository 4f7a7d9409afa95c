"""CPU check of the TMA-fed fp64 GEMM's shared-memory addressing (elemental_b200/csrc/gemm_f64_layout.hpp, used by
the experimental gemm_f64_tma.cu): tests/native/f64_tma_layout_emul.cpp fills a stage tile the way TMA with the
128-byte swizzle would, reads every fragment through the kernel's own offset functions, applies the m8n8k4
fragment semantics and the kernel's epilogue mapping, and must reproduce the tile product for NN / NT / TN / TT,
with every fragment load free of bank conflicts."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_f64_tma_layout_emulation(tmp_path):
    exe = str(tmp_path / "f64_tma_layout_emul")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(ROOT, "tests", "native", "f64_tma_layout_emul.cpp")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "F64_TMA_LAYOUT_PASS" in r.stdout, r.stdout + r.stderr
