"""Achieved HBM bandwidth of the BLAS-1 / data-movement kernels of libelb200 (C ABI), against the
measured copy bandwidth in MEASURED_PEAKS.json.  Algorithmic bytes per call as in DESIGN.md 3.3:
copy / transpose / pack / unpack 2*h*w*s, axpy and axpy-unpack 3*h*w*s, scale 2*h*w*s, fill h*w*s.
Usage: python tools/blas1_bw.py [--n 16384] [--out profiles/r01_blas1_bandwidth.json]"""
import argparse
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from elemental_b200 import _lib as L  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16384)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    peak = 6541.1
    mp = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(mp):
        try:
            peak = float(json.load(open(mp)).get("hbm_gbs", peak))
        except Exception:
            pass
    dev = torch.device("cuda:0")
    n = args.n
    rows = []
    for dt, sfx, ct in ((torch.float32, "f32", C.c_float), (torch.float64, "f64", C.c_double)):
        s = torch.tensor([], dtype=dt).element_size()
        X = torch.rand(n * n, device=dev, dtype=dt)
        Y = torch.rand(n * n, device=dev, dtype=dt)
        buf = torch.empty(n * n, device=dev, dtype=dt)
        lib = L.lib
        stride = 4
        segs = (L.Segment * stride)(*[L.Segment(0, 1, n, k, stride, n // stride, k * n * (n // stride)) for k in range(stride)])
        rsegs = (L.Segment * stride)(*[L.Segment(k, stride, n // stride, 0, 1, n, k * (n // stride) * n) for k in range(stride)])
        one = L.Segment(0, 1, n, 0, 1, n, 0)
        srcs = (C.c_void_p * 2)(X.data_ptr(), buf.data_ptr())
        cases = [
            ("copy2d", 2, lambda: getattr(lib, f"elb200_copy2d_{sfx}")(n, n, X.data_ptr(), 1, n, Y.data_ptr(), 1, n, None)),
            ("copy2d row-strided dst (ColStridedUnpack shape)", 2,
             lambda: getattr(lib, f"elb200_copy2d_{sfx}")(n // 2, n, X.data_ptr(), 1, n // 2, Y.data_ptr(), 2, n, None)),
            ("transpose", 2, lambda: getattr(lib, f"elb200_transpose_{sfx}")(n, n, X.data_ptr(), n, Y.data_ptr(), n, None)),
            ("axpy2d", 3, lambda: getattr(lib, f"elb200_axpy2d_{sfx}")(n, n, ct(0.5), X.data_ptr(), 1, n, Y.data_ptr(), 1, n, None)),
            ("axpy2d_transpose", 3, lambda: getattr(lib, f"elb200_axpy2d_transpose_{sfx}")(n, n, ct(0.5), X.data_ptr(), n, Y.data_ptr(), n, None)),
            ("scale2d", 2, lambda: getattr(lib, f"elb200_scale2d_{sfx}")(n, n, ct(1.0001), Y.data_ptr(), n, None)),
            ("fill2d", 1, lambda: getattr(lib, f"elb200_fill2d_{sfx}")(n, n, ct(0.0), Y.data_ptr(), n, None)),
            ("lattice PACK  column-interleaved x4 (RowStridedPack)", 2,
             lambda: getattr(lib, f"elb200_lattice_{sfx}")(L.PACK, stride, segs, X.data_ptr(), n, buf.data_ptr(), ct(0), None)),
            ("lattice UNPACK column-interleaved x4 (RowStridedUnpack)", 2,
             lambda: getattr(lib, f"elb200_lattice_{sfx}")(L.UNPACK, stride, segs, Y.data_ptr(), n, buf.data_ptr(), ct(0), None)),
            ("lattice UNPACK row-interleaved x4 (ColStridedUnpack)", 2,
             lambda: getattr(lib, f"elb200_lattice_{sfx}")(L.UNPACK, stride, rsegs, Y.data_ptr(), n, buf.data_ptr(), ct(0), None)),
            ("lattice UNPACK_AXPY (AxpyContract update)", 3,
             lambda: getattr(lib, f"elb200_lattice_{sfx}")(L.UNPACK_AXPY, stride, segs, Y.data_ptr(), n, buf.data_ptr(), ct(0.5), None)),
            ("peer_reduce_axpy, 2 sources (local)", 4,
             lambda: getattr(lib, f"elb200_peer_reduce_axpy_{sfx}")(2, srcs, 0, C.byref(one), Y.data_ptr(), n, ct(0.5), None)),
            # trapezoid kernel: algorithmic bytes count the entries INSIDE the lower triangle
            # (n(n+1)/2) for SCALE (2 passes) and AXPY (3), the entries OUTSIDE it (n(n-1)/2) for MAKE (1)
            ("trapezoid SCALE lower", 2, lambda: getattr(lib, f"elb200_trapezoid_{sfx}")(L.TRAP_SCALE, L.LOWER, 0, n, n, 0, 1, 0, 1, ct(1.0001), None, n, Y.data_ptr(), n, None)),
            ("trapezoid AXPY lower", 3, lambda: getattr(lib, f"elb200_trapezoid_{sfx}")(L.TRAP_AXPY, L.LOWER, 0, n, n, 0, 1, 0, 1, ct(0.5), X.data_ptr(), n, Y.data_ptr(), n, None)),
            ("trapezoid AXPY lower, element-cyclic 2x4 index map", 3,
             lambda: getattr(lib, f"elb200_trapezoid_{sfx}")(L.TRAP_AXPY, L.LOWER, 0, n, n, 1, 2, 3, 4, ct(0.5), X.data_ptr(), n, Y.data_ptr(), n, None)),
            ("trapezoid MAKE lower", 1, lambda: getattr(lib, f"elb200_trapezoid_{sfx}")(L.TRAP_MAKE, L.LOWER, 0, n, n, 0, 1, 0, 1, ct(0), None, n, Y.data_ptr(), n, None)),
        ]
        for name, passes, fn in cases:
            for _ in range(3):
                assert fn() == 0, name
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 10
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
            elems = (n // 2) * n if "row-strided dst" in name else n * n
            if name.startswith("trapezoid"):
                if "index map" in name:   # global (1 + 2i, 3 + 4j): inside when 1 + 2i >= 3 + 4j
                    elems = sum(max(0, n - max(0, (2 + 4 * j + 1) // 2)) for j in range(n))
                else:
                    elems = n * (n + 1) // 2
                if "MAKE" in name:
                    elems = n * n - elems
            gbs = passes * elems * s / (ms * 1e-3) / 1e9
            rows.append({"kernel": name, "dtype": sfx, "n": n, "ms": ms, "algorithmic_GBps": gbs, "frac_of_measured_hbm": gbs / peak})
            print(f"{sfx} {name:58s} {ms:8.3f} ms  {gbs:8.1f} GB/s  {gbs / peak:5.2f} of {peak:.0f}")
        del X, Y, buf
    if args.out:
        json.dump({"peak_hbm_GBps": peak, "note": "kernels launched on the legacy default stream (stream = NULL), CUDA-event timed, 10 repetitions after 3 warm-ups; matrices 1 GiB (fp32) / 2 GiB (fp64) >> 126 MB L2", "rows": rows}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
