"""Timeline MODEL of the engine's schedules, no GPU: the hostsim build of the engine (tests/hostsim) in dry-run mode
(HOSTSIM_DRY=1: no payload is computed or moved) with virtual time (HOSTSIM_VTIME=1): every stream carries a clock,
operations last bytes / rate or flops / rate with rates MEASURED on B200s (tests/hostsim/sim.hpp kRate*), and start
when their stream, their dependencies (events, arrival / ack flags, collective partners) and their resource (tensor
pipe, HBM kernels, PCIe up / down, NVLink egress / ingress) are ready.  cudaEventElapsedTime then returns modelled ms.

It answers "what does this schedule cost, and where does the time go" for bench-sized problems in seconds — the
numbers are a model to be CALIBRATED against measurements (see the table it prints), not measurements.

  python tools/model_timeline.py                     # the calibration table (known measurements next to the model)
  python tools/model_timeline.py gemm 2 4 32768 f32  # one scenario: El::Gemm SUMMA_C on a 2 x 4 grid
  python tools/model_timeline.py e2e 1 1 16384 f32 | c4 2 2 | c5 2 4 | syrk 2 4 32768 f32 | symm ... | trsm ...
"""
import ctypes as C
import os
import sys

os.environ.setdefault("HOSTSIM_VTIME", "1")
os.environ.setdefault("HOSTSIM_DRY", "1")
os.environ.setdefault("HOSTSIM_TIMEOUT", "300")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from tests.hostsim import sim_el as el  # noqa: E402

lib = el.lib
lib.elb200_grid_stream.restype = C.c_void_p
for f in ("cudaEventCreate", "cudaEventRecord", "cudaEventSynchronize", "cudaEventElapsedTime", "cudaEventDestroy"):
    getattr(lib, f).restype = C.c_int
lib.cudaEventRecord.argtypes = [C.c_void_p, C.c_void_p]
lib.cudaEventSynchronize.argtypes = [C.c_void_p]
lib.cudaEventElapsedTime.argtypes = [C.POINTER(C.c_float), C.c_void_p, C.c_void_p]


def timed(grid, fn, warm=1, reps=2) -> float:
    """modelled ms per call of fn on this rank (events on the grid's compute stream)"""
    st = lib.elb200_grid_stream(grid._h)
    for _ in range(warm):
        fn()
    grid.sync()
    e0, e1 = C.c_void_p(), C.c_void_p()
    lib.cudaEventCreate(C.byref(e0)); lib.cudaEventCreate(C.byref(e1))
    lib.cudaEventRecord(e0, st)
    for _ in range(reps):
        fn()
    lib.cudaEventRecord(e1, st)
    lib.cudaEventSynchronize(e1)
    grid.sync()
    ms = C.c_float()
    lib.cudaEventElapsedTime(C.byref(ms), e0, e1)
    return ms.value / reps


def dm(grid, U, V, dt, h, w):
    return el.DistMatrix(grid, U, V, dt).Resize(h, w)


def scen_gemm(r, c, n, dtype, alg=el.GEMM_SUMMA_C, m=None, k=None):
    dt = torch.float32 if dtype == "f32" else torch.float64
    m, k = m or n, k or n

    def body(g):
        A, B, Cm = dm(g, el.MC, el.MR, dt, m, k), dm(g, el.MC, el.MR, dt, k, n), dm(g, el.MC, el.MR, dt, m, n)
        return timed(g, lambda: el.Gemm(0, 0, 0.5, A, B, -0.5, Cm, alg))
    ms = max(el.run_grid(r, c, body))
    return ms, 2.0 * m * n * k / ms / 1e9


def scen_e2e(r, c, n, dtype):
    dt = torch.float32 if dtype == "f32" else torch.float64
    npdt = np.float32 if dtype == "f32" else np.float64

    def body(g):
        lh = -(-n // r)  # upper bound of the local height / width (host buffers are never touched in a dry run)
        lw = -(-n // c)
        bufs = [np.empty((lw, lh), dtype=npdt) for _ in range(3)]
        tens = [torch.from_numpy(b) for b in bufs]
        return timed(g, lambda: el.GemmHost(g, 0, 0, n, n, n, 0.5, tens[0], lh, tens[1], lh, -0.5, tens[2], lh), warm=1, reps=2)
    ms = max(el.run_grid(r, c, body))
    return ms, 2.0 * n ** 3 / ms / 1e9


def scen_c4(r, c, gbytes=1 << 30, dtype="f32"):
    dt = torch.float32 if dtype == "f32" else torch.float64
    esz = 4 if dtype == "f32" else 8
    n = int((gbytes / esz) ** 0.5) // 64 * 64

    def body(g):
        A = dm(g, el.MC, el.MR, dt, n, n)
        B1 = dm(g, el.MC, el.STAR, dt, n, n)
        B2 = dm(g, el.STAR, el.VR, dt, n, n)
        t1 = timed(g, lambda: el.Copy(A, B1))
        t2 = timed(g, lambda: el.Copy(A, B2))
        t3 = timed(g, lambda: el.AxpyContract(1e-3, B1, A))
        return (t1, t2, t3)
    res = el.run_grid(r, c, body)
    return tuple(max(x[i] for x in res) for i in range(3))


def scen_level3(r, c, n, dtype, what):
    """Syrk (C := a A A^T + b C, LOWER, n x n from n x n) / Symm (LEFT, LOWER) / Trsm (LEFT LOWER NORMAL, n x n right-hand sides)"""
    dt = torch.float32 if dtype == "f32" else torch.float64

    def body(g):
        A, B, Cm = dm(g, el.MC, el.MR, dt, n, n), dm(g, el.MC, el.MR, dt, n, n), dm(g, el.MC, el.MR, dt, n, n)
        if what == "syrk":
            return timed(g, lambda: el.Syrk(el.LOWER, 0, 0.5, A, -0.5, Cm))
        if what == "symm":
            return timed(g, lambda: el.Symm(el.LEFT, el.LOWER, 0.5, A, B, -0.5, Cm))
        return timed(g, lambda: el.Trsm(el.LEFT, el.LOWER, 0, el.NON_UNIT, 0.5, A, B))
    ms = max(el.run_grid(r, c, body))
    flops = {"syrk": 1.0, "symm": 2.0, "trsm": 1.0}[what] * n ** 3
    return ms, flops / ms / 1e9


def main():
    a = sys.argv[1:]
    if a and a[0] == "gemm":
        r, c, n, dt = int(a[1]), int(a[2]), int(a[3]), a[4]
        ms, tf = scen_gemm(r, c, n, dt)
        print(f"model: El::Gemm {dt} SUMMA_C n={n} on {r}x{c}: {ms:.2f} ms ({tf:.1f} TF/s)")
        return
    if a and a[0] == "e2e":
        r, c, n, dt = int(a[1]), int(a[2]), int(a[3]), a[4]
        ms, tf = scen_e2e(r, c, n, dt)
        print(f"model: host-operand El::Gemm {dt} n={n} on {r}x{c}: {ms:.2f} ms ({tf:.1f} TF/s)")
        return
    if a and a[0] == "c4":
        r, c = int(a[1]), int(a[2])
        t = scen_c4(r, c)
        print(f"model: C4 1 GiB fp32 on {r}x{c}: {t[0]:.2f} / {t[1]:.2f} / {t[2]:.2f} ms ([MC,MR]->[MC,*] / ->[*,VR] / AxpyContract)")
        return
    if a and a[0] in ("syrk", "symm", "trsm"):
        r, c, n, dt = int(a[1]), int(a[2]), int(a[3]), a[4]
        ms, tf = scen_level3(r, c, n, dt, a[0])
        print(f"model: El::{a[0].capitalize()} {dt} n={n} on {r}x{c}: {ms:.2f} ms ({tf:.1f} useful TF/s)")
        return
    if a and a[0] == "c5":
        r, c = int(a[1]), int(a[2])
        ms, tf = scen_gemm(r, c, 8192, "f32", alg=el.GEMM_DEFAULT, m=8192, k=262144)
        print(f"model: C5 fp32 m=n=8192 k=262144 GEMM_DEFAULT on {r}x{c}: {ms:.2f} ms ({tf:.1f} TF/s)")
        return
    # calibration table: measured numbers next to the model. Every row runs in a process
    # of its own (the engine reads its ELB200_* switches once), so the schedule that WAS measured can be modelled
    # next to the current default.
    import subprocess

    def sub(args, env=None):
        e = dict(os.environ, **(env or {}))
        out = subprocess.run([sys.executable, os.path.abspath(__file__)] + args, env=e, capture_output=True, text=True, timeout=900)
        line = [ln for ln in out.stdout.splitlines() if ln.startswith("model:")]
        return line[-1].split(": ", 2)[-1] if line else "FAILED " + out.stderr[-200:]

    rows = [
        ("C2 fp32 n=16384 1x1", "38.7 ms (227.5 TF/s)", ["gemm", "1", "1", "16384", "f32"], None),
        ("e2e fp32 n=16384 1x1, tapered tail, whole B slices (ELB200_HOST_BRANGES=1)", "64.4 ms (136.5 TF/s)", ["e2e", "1", "1", "16384", "f32"], {"ELB200_HOST_BRANGES": "1"}),
        ("e2e fp32 n=16384 1x1, tapered tail + B in 4 k ranges (default)", "62.4 ms (140.9 TF/s)", ["e2e", "1", "1", "16384", "f32"], None),
        ("fp32 n=32768 2x4", "40.9 ms (1720.9 TF/s)", ["gemm", "2", "4", "32768", "f32"], None),
        ("C3 fp64 n=32768 2x4", "292 ms (240.7 TF/s)", ["gemm", "2", "4", "32768", "f64"], None),
        ("C4 1 GiB 2x2, gather in one piece (round-1 default)", "0.54 / 0.55 / 0.61 ms", ["c4", "2", "2"], {"ELB200_AG_SPLIT_BYTES": "0"}),
        ("C4 1 GiB 2x2, split gather with rotation (default since round 2)", "0.475 / 0.55 / 0.61 ms", ["c4", "2", "2"], None),
        ("C4 1 GiB 2x4, gather in one piece", "1.28 / 0.36 / 0.82 ms", ["c4", "2", "4"], {"ELB200_AG_SPLIT_BYTES": "0"}),
        ("C4 1 GiB 2x4, split gather with rotation (default)", "see profiles/r02_dist_sweep_8gpu.json", ["c4", "2", "4"], None),
        ("C5 fp32 k=262144 2x4, k-range pipeline (default)", "21.7 ms (1624 TF/s)", ["c5", "2", "4"], None),
    ]
    w = max(len(r[0]) for r in rows)
    print(f"{'scenario':{w}s} | measured on B200 | model")
    for label, meas, args, env in rows:
        print(f"{label:{w}s} | {meas} | {sub(args, env)}", flush=True)


if __name__ == "__main__":
    main()
