"""Builds tests/hostsim/_build/libelb200_hostsim.so: libelb200's host ENGINE sources (unchanged, from
elemental_b200/csrc) linked against the hostsim test doubles instead of the CUDA runtime, NCCL and
the device kernels. Test infrastructure only — see tests/hostsim/sim.hpp."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "elemental_b200", "csrc")
# HOSTSIM_SANITIZE=address|thread|undefined builds an instrumented copy into _build_<name>/ (run the python
# worker under LD_PRELOAD of the matching runtime, e.g. $(gcc -print-file-name=libtsan.so))
SAN = os.environ.get("HOSTSIM_SANITIZE", "")
OUT = os.path.join(HERE, "_build" + ("_" + SAN if SAN else ""))
LIB = os.path.join(OUT, "libelb200_hostsim.so")
ENGINE = ["engine_core.cpp", "engine_gemm.cpp", "engine_redist.cpp", "engine_level3.cpp", "peer.cpp", "plan.cpp", "capi_engine.cpp", "capi_complex.cpp", "comm.cpp"]
DOUBLES = ["sim_cuda.cpp", "sim_nccl.cpp", "sim_ops.cpp"]
CUDA_INC = os.environ.get("CUDA_HOME", "/usr/local/cuda") + "/include"
FLAGS = ["-std=c++17", "-O2", "-g", "-fPIC", "-pthread", "-Wall", "-Wno-unused-function", "-DELB200_HOSTSIM",
         "-I" + CUDA_INC, "-I/usr/include", "-I" + HERE]
if SAN:
    FLAGS += ["-fsanitize=" + SAN, "-fno-omit-frame-pointer", "-O1"]


def _newer(src: str, deps, obj: str) -> bool:
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src] + deps)


def build(verbose: bool = False) -> str:
    os.makedirs(OUT, exist_ok=True)
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".hpp", ".cuh"))]
    hdrs += [os.path.join(ROOT, "include", "elb200.h"), os.path.join(HERE, "sim.hpp"), os.path.abspath(__file__)]
    jobs, objs = [], []
    for name, base in [(f, CSRC) for f in ENGINE if os.path.exists(os.path.join(CSRC, f))] + [(f, HERE) for f in DOUBLES]:
        src, obj = os.path.join(base, name), os.path.join(OUT, name + ".o")
        objs.append(obj)
        if _newer(src, hdrs, obj):
            cmd = ["g++"] + FLAGS + ["-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd))
            jobs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in jobs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"hostsim: compiling {src} failed:\n{out[-6000:]}")
    if jobs or not os.path.exists(LIB):
        # -Bsymbolic: the engine's calls to cudaMalloc / ncclAllGather / elb200_gemm_f32 ... bind to the doubles
        # inside this library even if the real libcudart / libnccl / libelb200 are loaded in the process
        cmd = ["g++", "-shared", "-pthread", "-Wl,-Bsymbolic", "-o", LIB] + objs + (["-fsanitize=" + SAN] if SAN else [])
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"hostsim: link failed:\n{r.stdout[-6000:]}")
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
