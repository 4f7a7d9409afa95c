"""ORACLE — CPU restatement of the reference's El::Gemm path. TEST INFRASTRUCTURE ONLY.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs; the product (elemental_b200/) never imports it. See oracle/README.md.
"""
from . import dist, gemm  # noqa: F401
