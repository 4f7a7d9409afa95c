#!/usr/bin/env python
"""Static evidence that needs no GPU: per-kernel resource usage (ptxas -v) and a census of the SASS
mnemonics that show which hardware path each object uses (UTCHMMA = tcgen05.mma, UTMALDG = TMA load,
SYNCS = mbarrier, DMMA = fp64 tensor pipe, LDGSTS = cp.async, LDG/STG.E.128 = 128-bit global access).

    python tools/sass_census.py > profiles/rNN_sass_census.txt

Recompiles elemental_b200/csrc/*.cu into a scratch directory with the Makefile's flags plus -Xptxas -v."""
from __future__ import annotations

import collections
import glob
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "-Xptxas", "-v", "-I/usr/include"]
MNEMONICS = re.compile(r"\b(UTCHMMA|UTCQMMA|UTMALDG|UTMASTG|UTMAPF|UTCBAR|UTCATOMSWS|SYNCS|LDTM|STTM|DMMA|HMMA|LDGSTS|LDSM|UBLKCP|"
                       r"LDG\.E\.128|STG\.E\.128|LDG\.E\.64|STG\.E\.64|SHFL|REDG|ATOMG|ATOMS|MEMBAR|FENCE)\b[A-Z0-9_.]*")
DROP = re.compile(r"\.(U32|SYS|STRONG|CONSTANT|GPU|SM|CTA|64|32)\b")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    short = []
    for n in out:
        n = re.sub(r"\(anonymous namespace\)::", "", n)
        n = re.sub(r"^void ", "", n)
        n = re.sub(r"\(.*$", "", n)  # drop the argument list
        short.append(n)
    return short


def main() -> int:
    srcs = sorted(glob.glob(os.path.join(ROOT, "elemental_b200", "csrc", "*.cu")))
    tmp = tempfile.mkdtemp(prefix="sass_census_")
    procs = []
    for s in srcs:
        o = os.path.join(tmp, os.path.basename(s)[:-3] + ".o")
        procs.append((s, o, subprocess.Popen(["nvcc", *FLAGS, "-c", s, "-o", o], stderr=subprocess.PIPE, text=True)))
    print("# nvcc " + " ".join(FLAGS))
    print(subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout.strip().splitlines()[-2])
    for s, o, p in procs:
        _, err = p.communicate()
        if p.returncode != 0:
            print(f"\n== {os.path.basename(s)}: nvcc failed\n{err[-2000:]}")
            return 1
        print(f"\n== {os.path.relpath(s, ROOT)}")
        # ptxas -v: "Compiling entry function 'X'" / "Function properties" / "N bytes stack frame, ... spill" / "Used N registers, ..."
        entries, cur = [], None
        for ln in err.splitlines():
            m = re.search(r"Compiling entry function '([^']+)'", ln)
            if m:
                cur = {"name": m.group(1), "spill": "", "used": ""}
                entries.append(cur)
            elif cur is not None and "bytes stack frame" in ln:
                cur["spill"] = ln.strip()
            elif cur is not None and "Used" in ln and "registers" in ln:
                cur["used"] = ln.split(":", 1)[1].strip()
        names = demangle([e["name"] for e in entries]) if entries else []
        for e, n in zip(entries, names):
            spill = e["spill"]
            spill = "no spills" if re.search(r"0 bytes spill stores, 0 bytes spill loads", spill) and spill.startswith("0 bytes stack") else spill
            print(f"  {n}: {e['used']}; {spill}")
        sass = subprocess.run(["cuobjdump", "-sass", o], capture_output=True, text=True).stdout
        cnt = collections.Counter(DROP.sub("", m.group(0)) for m in MNEMONICS.finditer(sass))
        if cnt:
            top = cnt.most_common(16)
            # the mnemonics that prove TMA / tensor-core use are listed even when they are rare
            top += [(k, v) for k, v in sorted(cnt.items()) if (k, v) not in top and any(t in k for t in ("UTMALDG", "UTCHMMA", "UTCBAR", "DMMA", "LDGSTS"))]
            print("  SASS census (all kernels of the object): " + ", ".join(f"{k} x{v}" for k, v in top))
    return 0


if __name__ == "__main__":
    sys.exit(main())
